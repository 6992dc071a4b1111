"""Scene ingest (N3): er_load_obj (OBJ/MTL text -> triangles: parse, triangulation, z flip, optional
face-weighted normals, MikkTSpace tangents; elevenrender_b200/host/er_objloader.hpp, er_mikktspace.hpp)
against the reference's own loader (ObjLoader::loadObjsRapid stream path + the vendored MikkTSpace) compiled
into the oracle. Every word of every triangle must match bit for bit. CPU only."""
import numpy as np
import pytest

import common
from elevenrender_b200 import capi, scenes


def grid_obj(nx, ny, quads, seed, shared=True, two_objects=False, flip_uv=False):
    """a displaced grid with shared vertices, smooth normals and a real uv parametrisation"""
    rng = np.random.default_rng(seed)
    xs, ys = np.linspace(-1, 1, nx + 1), np.linspace(-1, 1, ny + 1)
    X, Y = np.meshgrid(xs, ys, indexing="xy")
    Z = 0.3 * np.sin(3 * X + 0.5) * np.cos(2 * Y) + 0.05 * rng.standard_normal(X.shape)
    N = np.stack([-np.gradient(Z, axis=1) * nx / 2, -np.gradient(Z, axis=0) * ny / 2, np.ones_like(Z)], -1)
    N /= np.linalg.norm(N, axis=-1, keepdims=True)
    U, V = (X + 1) / 2, (Y + 1) / 2
    if flip_uv:
        U = 1 - U  # mirrored parametrisation: orientation-reversing triangles
    lines = ["mtllib m.mtl", "o grid"]
    idx = lambda i, j: j * (nx + 1) + i + 1
    for j in range(ny + 1):
        for i in range(nx + 1):
            lines.append("v %.9g %.9g %.9g" % (X[j, i], Y[j, i], Z[j, i]))
            lines.append("vt %.9g %.9g" % (U[j, i], V[j, i]))
            lines.append("vn %.9g %.9g %.9g" % tuple(N[j, i]))
    for j in range(ny):
        if two_objects and j == ny // 2:
            lines.append("g upper")
        lines.append("usemtl %s" % ("a" if j % 2 else "b"))
        for i in range(nx):
            a, b, c, d = idx(i, j), idx(i + 1, j), idx(i + 1, j + 1), idx(i, j + 1)
            f = lambda k: "%d/%d/%d" % (k, k, k)
            if quads:
                lines.append("f " + " ".join(f(k) for k in (a, b, c, d)))
            else:
                lines.append("f " + " ".join(f(k) for k in (a, b, c)))
                lines.append("f " + " ".join(f(k) for k in (a, c, d)))
    return "\n".join(lines) + "\n", "newmtl a\nKd 1 0 0\nnewmtl b\nKd 0 1 0\n"


def sphere_obj(n_lat=9, n_lon=14):
    """lat-long sphere: poles (degenerate uv strips), a uv seam, smooth normals"""
    lines = ["o ball"]
    vid = {}
    for a in range(n_lat + 1):
        th = np.pi * a / n_lat
        for b in range(n_lon + 1):
            ph = 2 * np.pi * b / n_lon
            p = (np.sin(th) * np.cos(ph), np.cos(th), np.sin(th) * np.sin(ph))
            lines.append("v %.9g %.9g %.9g" % p)
            lines.append("vt %.9g %.9g" % (b / n_lon, a / n_lat))
            lines.append("vn %.9g %.9g %.9g" % p)
            vid[a, b] = len(vid) + 1
    for a in range(n_lat):
        for b in range(n_lon):
            q = [vid[a, b], vid[a, b + 1], vid[a + 1, b + 1], vid[a + 1, b]]
            lines.append("f " + " ".join("%d/%d/%d" % (k, k, k) for k in q))
    return "\n".join(lines) + "\n", ""


CASES = {
    "cornell": lambda: scenes.cornell_obj_text(),
    "grid_tris": lambda: grid_obj(7, 5, False, 1),
    "grid_quads": lambda: grid_obj(6, 6, True, 2),
    "grid_two_objects": lambda: grid_obj(5, 6, True, 3, two_objects=True),
    "grid_mirrored_uv": lambda: grid_obj(5, 4, False, 4, flip_uv=True),
    "sphere": lambda: sphere_obj(),
}


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("recompute", [False, True])
def test_ingest_matches_reference_loader(ref, name, recompute):
    obj, mtl = CASES[name]()
    rs = ref.RefScene(8, 8)
    with ref._quiet():
        n = rs.load_obj(obj, mtl, recompute)
    want = rs.get_tris()
    rs.close()
    got, names = capi.load_obj(obj, mtl, recompute)
    assert got.shape[0] == n == want.shape[0]
    labels = [("vertices", 0, 9), ("uv", 9, 18), ("normals", 18, 27), ("tangents", 27, 36), ("sign", 36, 37), ("objectID", 37, 38)]
    for what, a, b in labels:
        same = got[:, a:b].view(np.uint32) == want[:, a:b].view(np.uint32)
        # +-0 are the same value to every consumer
        same |= (got[:, a:b] == 0) & (want[:, a:b] == 0)
        bad = np.nonzero(~same.all(axis=1))[0]
        assert bad.size == 0, f"{name} recompute={recompute}: {what} differ in {bad.size} of {n} triangles, first {bad[:5]}: " \
                              f"{got[bad[0], a:b]} vs {want[bad[0], a:b]}"


def test_ingest_material_names(ref):
    obj, mtl = grid_obj(3, 4, True, 5)
    tris, names = capi.load_obj(obj, mtl)
    assert len(names) == tris.shape[0] == 24
    assert set(names) == {"a", "b"} and names[0] == "b" and names[6] == "a"
