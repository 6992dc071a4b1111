"""Procedural scenes of the BASELINE.json configs, generated with numpy.

Every generator returns flat ErTri words ((n, 39) float32: vertices9 uv9 normals9 tangents9 sign
objectID materialID, include/er_cuda.h) so the same arrays feed the CUDA path (Scene.addMeshObject)
and the reference oracle (tests/refharness.py).  Rules that keep the reference well-behaved:
camera and ray origins stay off the world origin (empty BVH nodes are boxes at 0),
vertex normals face the side rays arrive from (F10: the shading normal is not face-forwarded),
textures are LINEAR/NO_FILTER (F11).  Seed 0x11E7 unless stated.
"""
import numpy as np

from .capi import TRI_WORDS

SEED = 0x11E7


def pack_tris(v, n, uv=None, t=None, sign=1.0, object_id=0, material_id=0):
    """v, n: (m,3,3) vertices / vertex normals; uv: (m,3,2) or None (-> (0,0),(1,0),(0,1), which makes
    Hit.tu/tv equal the barycentrics u,v exactly, Tri.h:76); t: (m,3,3) tangents or None."""
    v = np.asarray(v, dtype=np.float32)
    m = v.shape[0]
    w = np.zeros((m, TRI_WORDS), dtype=np.float32)
    w[:, 0:9] = v.reshape(m, 9)
    uvw = np.zeros((m, 3, 3), dtype=np.float32)
    if uv is None:
        uvw[:, 1, 0] = 1.0
        uvw[:, 2, 1] = 1.0
    else:
        uvw[:, :, 0:2] = np.asarray(uv, dtype=np.float32)
    w[:, 9:18] = uvw.reshape(m, 9)
    n = np.asarray(n, dtype=np.float32)
    w[:, 18:27] = n.reshape(m, 9)
    if t is None:
        # any unit vector orthogonal to the normal
        a = np.where(np.abs(n[..., 0:1]) < 0.9, np.float32([1, 0, 0]), np.float32([0, 1, 0]))
        t = np.cross(a, n)
        t = t / np.maximum(np.linalg.norm(t, axis=-1, keepdims=True), 1e-20)
    w[:, 27:36] = np.asarray(t, dtype=np.float32).reshape(m, 9)
    w[:, 36] = sign
    w[:, 37].view(np.int32)[:] = object_id
    w[:, 38].view(np.int32)[:] = np.asarray(material_id, dtype=np.int32)
    return w


def _quad(p0, p1, p2, p3, normal):
    """two triangles (p0,p1,p2),(p0,p2,p3) sharing one flat normal"""
    v = np.float32([[p0, p1, p2], [p0, p2, p3]])
    n = np.broadcast_to(np.float32(normal), (2, 3, 3)).copy()
    uv = np.float32([[[0, 0], [1, 0], [1, 1]], [[0, 0], [1, 1], [0, 1]]])
    return v, n, uv


def _box(lo, hi, outward=True):
    lo, hi = np.float32(lo), np.float32(hi)
    x0, y0, z0 = lo
    x1, y1, z1 = hi
    s = 1.0 if outward else -1.0
    faces = [
        ((x0, y0, z0), (x0, y1, z0), (x1, y1, z0), (x1, y0, z0), (0, 0, -s)),
        ((x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1), (0, 0, s)),
        ((x0, y0, z0), (x0, y0, z1), (x0, y1, z1), (x0, y1, z0), (-s, 0, 0)),
        ((x1, y0, z0), (x1, y1, z0), (x1, y1, z1), (x1, y0, z1), (s, 0, 0)),
        ((x0, y0, z0), (x1, y0, z0), (x1, y0, z1), (x0, y0, z1), (0, -s, 0)),
        ((x0, y1, z0), (x0, y1, z1), (x1, y1, z1), (x1, y1, z0), (0, s, 0)),
    ]
    return faces


def cornell_box():
    """C1: 5 walls + 2 boxes + one emissive ceiling quad = 36 triangles in the cube [-1,1]x[0,2]x[-1,1];
    camera at (0,1,-3.4) looking down +z.  Materials: 0 default (grey), 1 red, 2 green, 3 light.
    Returns (tri_words, materials) with materials as keyword dicts for host.material_words."""
    parts = []

    def add(faces, mat):
        for f in faces:
            v, n, uv = _quad(*f)
            parts.append(pack_tris(v, n, uv, material_id=mat))

    room = _box((-1, 0, -1), (1, 2, 1), outward=False)
    # room faces: 0 front (z=-1, open: skipped), 1 back, 2 left, 3 right, 4 floor, 5 ceiling
    add([room[1], room[4], room[5]], 0)
    add([room[2]], 1)
    add([room[3]], 2)
    add(_box((-0.65, 0.0, 0.0), (-0.1, 1.2, 0.55)), 0)
    add(_box((0.15, 0.0, -0.55), (0.7, 0.6, 0.0)), 0)
    add([((-0.3, 1.995, -0.3), (0.3, 1.995, -0.3), (0.3, 1.995, 0.3), (-0.3, 1.995, 0.3), (0, -1, 0))], 3)
    tris = np.concatenate(parts, axis=0)
    materials = [dict(albedo=(0.63, 0.065, 0.05)), dict(albedo=(0.14, 0.45, 0.091)),
                 dict(albedo=(0.78, 0.78, 0.78), emission=(8.0, 8.0, 8.0))]
    return tris, materials


def cornell_obj_text():
    """The same box as Wavefront OBJ + MTL text, for the reference's ObjLoader stream path
    (ObjLoader.cpp:154-164).  Normals are given per face; uvs per corner.  The loader negates z of
    positions and normals (ObjLoader.cpp:69-147), so the text carries -z: the loaded scene is cornell_box()."""
    tris, _ = cornell_box()
    tris = tris.copy()
    for k in range(3):
        tris[:, 3 * k + 2] *= -1.0
        tris[:, 18 + 3 * k + 2] *= -1.0
    mat_names = ["default", "red", "green", "light"]
    lines, vi = ["mtllib cornell.mtl", "o cornell"], 1
    cur = None
    for w in tris:
        mid = int(w[38:39].view(np.int32)[0])
        if mid != cur:
            lines.append(f"usemtl {mat_names[mid]}")
            cur = mid
        for k in range(3):
            lines.append("v %.9g %.9g %.9g" % tuple(w[3 * k:3 * k + 3]))
            lines.append("vt %.9g %.9g" % tuple(w[9 + 3 * k:9 + 3 * k + 2]))
            lines.append("vn %.9g %.9g %.9g" % tuple(w[18 + 3 * k:18 + 3 * k + 3]))
        lines.append("f %d/%d/%d %d/%d/%d %d/%d/%d" % (vi, vi, vi, vi + 1, vi + 1, vi + 1, vi + 2, vi + 2, vi + 2))
        vi += 3
    mtl = "\n".join(f"newmtl {n}\nKd 0.5 0.5 0.5" for n in mat_names) + "\n"
    return "\n".join(lines) + "\n", mtl


def cornell_camera():
    return dict(position=(0.0, 1.0, -3.4), rotation=(0.0, 0.0, 0.0))


def displaced_grid(nx=354, ny=353, x_half=1.6, y_half=1.0, z0=3.0, amp=0.2, fx=7.0, fy=5.0, unit_uv=True,
                   material_cells=0, object_id=0):
    """C2: z = z0 + amp*sin(fx*x)*cos(fy*y) over [-x_half,x_half]x[-y_half,y_half], nx*ny cells, two
    triangles per cell (354x353 -> 249 924), analytic vertex normals facing -z (towards the camera).
    unit_uv=True gives every triangle uv (0,0),(1,0),(0,1) (barycentric read-back); otherwise uv is
    the surface parametrisation in [0,1]^2.  material_cells=k assigns material 1 + (cell checker of
    k x k blocks) % 16 (C4)."""
    xs = np.linspace(-x_half, x_half, nx + 1, dtype=np.float64)
    ys = np.linspace(-y_half, y_half, ny + 1, dtype=np.float64)
    X, Y = np.meshgrid(xs, ys, indexing="xy")  # (ny+1, nx+1)
    Z = z0 + amp * np.sin(fx * X) * np.cos(fy * Y)
    dzdx = amp * fx * np.cos(fx * X) * np.cos(fy * Y)
    dzdy = -amp * fy * np.sin(fx * X) * np.sin(fy * Y)
    N = np.stack([dzdx, dzdy, -np.ones_like(Z)], axis=-1)
    N /= np.linalg.norm(N, axis=-1, keepdims=True)
    T = np.stack([np.ones_like(Z), np.zeros_like(Z), dzdx], axis=-1)
    T /= np.linalg.norm(T, axis=-1, keepdims=True)
    P = np.stack([X, Y, Z], axis=-1)
    UV = np.stack([(X + x_half) / (2 * x_half), (Y + y_half) / (2 * y_half)], axis=-1)

    def corners(A):
        return A[:-1, :-1], A[:-1, 1:], A[1:, 1:], A[1:, :-1]  # p00 p10 p11 p01

    def tri_pair(A):
        a, b, c, d = corners(A)
        t1 = np.stack([a, b, c], axis=2)  # (ny, nx, 3, k)
        t2 = np.stack([a, c, d], axis=2)
        return np.stack([t1, t2], axis=2).reshape(-1, 3, A.shape[-1])

    v, n, t = tri_pair(P), tri_pair(N), tri_pair(T)
    uv = None if unit_uv else tri_pair(UV)
    mat = 0
    if material_cells:
        cy, cx = np.meshgrid(np.arange(ny) // material_cells, np.arange(nx) // material_cells, indexing="ij")
        m = 1 + (cx * 5 + cy * 3) % 16
        mat = np.repeat(m.reshape(-1), 2)
    return pack_tris(v, n, uv, t, material_id=mat, object_id=object_id)


def grid_camera():
    return dict(position=(0.013, 0.017, -0.9), rotation=(0.0, 0.0, 0.0))


def _hash01(a):
    """integer hash -> [0,1) (Wang-style), vectorised"""
    a = a.astype(np.uint64)
    a = (a ^ 61) ^ (a >> 16)
    a = (a * 9) & 0xFFFFFFFF
    a = a ^ (a >> 4)
    a = (a * 0x27D4EB2D) & 0xFFFFFFFF
    a = a ^ (a >> 15)
    return (a & 0xFFFFFF).astype(np.float64) / float(1 << 24)


def sphere_field(side=25, lon=125, lat=64, spacing=0.15, radius=0.06, z0=4.0, noise=0.12):
    """C3: side*side spheres on a plane at z0, each a lon x lat lat-long mesh with two triangles per
    cell (25x25 x 125x64x2 = 10 000 000), radially displaced by hash noise; vertex normals are the
    undisplaced outward directions.  Degenerate pole triangles are kept (they never pass the
    determinant test), as an OBJ exporter would."""
    th = np.linspace(0.0, np.pi, lat + 1)            # polar
    ph = np.linspace(0.0, 2 * np.pi, lon + 1)        # azimuth
    TH, PH = np.meshgrid(th, ph, indexing="ij")      # (lat+1, lon+1)
    D = np.stack([np.sin(TH) * np.cos(PH), np.cos(TH), np.sin(TH) * np.sin(PH)], axis=-1)
    Tg = np.stack([-np.sin(PH), np.zeros_like(PH), np.cos(PH)], axis=-1)
    gi, gj = np.meshgrid(np.arange(lat + 1), np.arange(lon + 1) % lon, indexing="ij")  # seam shares noise
    out = []
    half = (side - 1) * 0.5
    for sy in range(side):
        rows = []
        for sx in range(side):
            sid = sy * side + sx
            h = _hash01(gi * 131 + gj * 7919 + sid * 104729 + SEED)
            h[0, :] = h[0, 0]
            h[-1, :] = h[-1, 0]
            r = radius * (1.0 + noise * (h - 0.5))
            c = np.array([(sx - half) * spacing, (sy - half) * spacing, z0])
            P = c + D * r[..., None]

            def pair(A):
                a, b, cc, d = A[:-1, :-1], A[:-1, 1:], A[1:, 1:], A[1:, :-1]
                t1 = np.stack([a, b, cc], axis=2)
                t2 = np.stack([a, cc, d], axis=2)
                return np.stack([t1, t2], axis=2).reshape(-1, 3, 3)

            rows.append(pack_tris(pair(P), pair(D), None, pair(Tg), object_id=sid))
        out.append(np.concatenate(rows, axis=0))
    return np.concatenate(out, axis=0)


def sphere_camera():
    return dict(position=(0.011, 0.023, -0.2), rotation=(0.0, 0.0, 0.0))


def sky_hdri(width=2048, height=1024, sun_peak=50.0, seed=SEED):
    """Synthetic lat-long environment: 0.05 ambient + vertical sky gradient + gaussian sun.
    Returns (height*width*3,) float32 texels (LINEAR)."""
    v = (np.arange(height, dtype=np.float64) + 0.5) / height
    u = (np.arange(width, dtype=np.float64) + 0.5) / width
    U, V = np.meshgrid(u, v, indexing="xy")
    sky = 0.05 + 0.6 * (1.0 - V)[..., None] * np.array([0.45, 0.65, 1.0]) + 0.15 * V[..., None] * np.array([0.9, 0.8, 0.7])
    su, sv, sig = 0.31, 0.27, 0.012
    du = np.minimum(np.abs(U - su), 1 - np.abs(U - su))
    sun = sun_peak * np.exp(-(du * du + (V - sv) ** 2) / (2 * sig * sig))
    img = sky + sun[..., None] * np.array([1.0, 0.93, 0.8])
    return img.astype(np.float32).reshape(-1)


def material_maps(size=1024, seed=SEED):
    """C4: procedural albedo (3ch), roughness (1ch), metallic (1ch) and tangent-space normal (3ch)
    maps, all LINEAR / NO_FILTER."""
    a = (np.arange(size, dtype=np.float64) + 0.5) / size
    U, V = np.meshgrid(a, a, indexing="xy")
    checker = ((np.floor(U * 16) + np.floor(V * 16)) % 2)
    albedo = np.stack([0.25 + 0.6 * checker, 0.3 + 0.4 * U, 0.35 + 0.4 * V], axis=-1)
    rough = 0.15 + 0.8 * (0.5 + 0.5 * np.sin(40 * U) * np.sin(40 * V))
    metal = (np.sin(25 * (U + V)) > 0.3).astype(np.float64) * 0.9
    nx, ny = 0.25 * np.cos(60 * U), 0.25 * np.cos(60 * V)
    nz = np.sqrt(np.maximum(1 - nx * nx - ny * ny, 0.0))
    normal = np.stack([nx, ny, nz], axis=-1) * 0.5 + 0.5
    f = lambda x: x.astype(np.float32).reshape(-1)
    return dict(albedo=f(albedo), roughness=f(rough), metallic=f(metal), normal=f(normal), size=size)


def random_rays(n, origin_box=((-2, -1.5, -1.0), (2, 1.5, 0.5)), target_box=((-1.6, -1.0, 2.7), (1.6, 1.0, 3.3)),
                seed=SEED):
    """n rays from random origins towards random targets (for er_trace parity): (n,6) float32"""
    rng = np.random.default_rng(seed)
    o = rng.uniform(origin_box[0], origin_box[1], size=(n, 3))
    t = rng.uniform(target_box[0], target_box[1], size=(n, 3))
    return np.concatenate([o, t - o], axis=1).astype(np.float32)
