"""T1 (-m gpu): device functions against the reference's own functions (oracle) on the same inputs.
Integer / comparison paths bit-exact; transcendental paths within a few ulp (CUDA libm vs glibc)."""
import os

import numpy as np
import pytest

import common
from elevenrender_b200 import host

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pair(gpu, ref):
    spec = common.make_spec("textured")
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], rng_order=host.ER_RNG_RTL))
    rm.upload(common.to_ours(spec))
    rs = common.to_ref(spec)
    yield spec, rm, rs
    rm.close()
    rs.close()


def test_rng_stream(gpu, ref):
    import ctypes as C
    from elevenrender_b200.capi import check, lib, ptr
    rm = host.RenderingManager()
    for idx in (0, 1, 77, 2073599, 8294399):
        st = np.zeros(64, np.uint32)
        v = np.zeros(64, np.float32)
        check(lib().er_test_rng(rm.ctx, idx, 64, ptr(st), ptr(v)))
        rst, rv = ref.rng_stream(idx, 64)
        assert np.array_equal(st, rst) and np.array_equal(v.view(np.uint32), rv.view(np.uint32))
    g = np.load(os.path.join(common.GOLDEN_DIR, "kat.npz"))
    st = np.zeros(16, np.uint32)
    v = np.zeros(16, np.float32)
    check(lib().er_test_rng(rm.ctx, 12345, 16, ptr(st), ptr(v)))
    assert np.array_equal(st, g["rng_state_12345"])
    rm.close()


def test_libm_bits(gpu, ref):
    """the device's sin/cos/acos/atan2/pow/log (er_libm.cuh) return the bits of the reference's libm"""
    import ctypes as C
    from elevenrender_b200.capi import check, lib, ptr
    rm = host.RenderingManager()
    rng = np.random.default_rng(11)
    n = 2_000_000
    ang = rng.uniform(-2 * np.pi, 4 * np.pi, n).astype(np.float32)
    unit = rng.uniform(-1, 1, n).astype(np.float32)
    unit[:8] = [1, -1, 0, 0.5, -0.5, 1e-9, 0.99999994, -0.99999994]
    y2, x2 = rng.normal(size=n).astype(np.float32), rng.normal(size=n).astype(np.float32)
    y2[:6], x2[:6] = [0, 0, 1, -1, 1e-30, 0], [1, -1, 0, 0, 1, 0]
    pos = np.abs(rng.normal(size=n)).astype(np.float32) + np.float32(1e-12)
    base = rng.uniform(0, 1, n).astype(np.float32)
    base[:3] = [0, 1, 0.5]
    e22 = np.full(n, 2.2, np.float32)
    eany = rng.uniform(0.1, 6, n).astype(np.float32)
    cases = [("sin", 0, ang, None, 1e-7), ("cos", 1, ang, None, 1e-7), ("acos", 2, unit, None, 0.0),
             ("atan2", 3, y2, x2, 0.0), ("pow(x,2.2)", 4, base, e22, 1e-7), ("pow", 4, pos, eany, 1e-7), ("log", 5, pos, None, 0.0)]
    for name, fn, a, b, budget in cases:
        out = np.zeros(n, np.float32)
        check(lib().er_test_libm(rm.ctx, fn, ptr(a), ptr(b) if b is not None else None, C.c_size_t(n), ptr(out)))
        want = ref.libm(fn, a, b)
        bad = int((out.view(np.uint32) != want.view(np.uint32)).sum())
        assert bad <= budget * n + (2 if budget else 0), f"{name}: {bad} of {n} differ from the host libm"
    rm.close()


def test_texture_lookups(pair):
    import ctypes as C
    from elevenrender_b200.capi import check, lib, ptr
    spec, rm, rs = pair
    rng = np.random.default_rng(3)
    uv = rng.uniform(-2.5, 2.5, size=(4000, 2)).astype(np.float32)
    uv[:16] = [[0, 0], [1, 1], [1, 0], [0, 1], [-0.2, 0.3], [1.0000001, 0.5], [-1, -1], [0.999999, 0.999999],
               [2, 2], [-0.0, 0.0], [0.5, -0.5], [1e-8, 1e-8], [3.5, -3.5], [0.25, 0.75], [-2.5, 2.5], [1.5, 1.5]]
    px = rng.integers(-200, 200, size=(2000, 2)).astype(np.float32)
    for tex in (-1, 0, 1, 2, 3):
        for mode, ab in ((0, px), (1, uv), (2, uv), (3, uv)):
            out = np.zeros((ab.shape[0], 3), np.float32)
            check(lib().er_test_texture(rm.ctx, tex, mode, ptr(ab), C.c_size_t(ab.shape[0]), ptr(out)))
            want = rs.texture_lookup(tex, mode, ab)
            assert np.array_equal(out.view(np.uint32), want.view(np.uint32)), f"tex {tex} mode {mode}"


def test_spherical_mapping(pair):
    import ctypes as C
    from elevenrender_b200.capi import check, lib, ptr
    spec, rm, rs = pair
    rng = np.random.default_rng(5)
    d = rng.normal(size=(20000, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    d[:6] = [[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]]
    out = np.zeros((d.shape[0], 2), np.float32)
    check(lib().er_test_spherical(rm.ctx, ptr(d), C.c_size_t(d.shape[0]), ptr(out)))
    want = ref_sph = __import__("refharness").spherical_mapping(d)
    # acosf/atan2f differ by <= 2 ulp between CUDA and glibc; limitUV may wrap u at the seam
    du = np.abs(out - want)
    du[:, 0] = np.minimum(du[:, 0], np.abs(du[:, 0] - 1.0))
    assert du.max() < 1e-6


def test_hdri_sample_chain(pair):
    import ctypes as C
    from elevenrender_b200.capi import check, lib, ptr
    spec, rm, rs = pair
    rng = np.random.default_rng(9)
    r = rng.random(20000).astype(np.float32)
    r[:4] = [0.0, 1.0, 0.5, 1e-9]
    out = np.zeros((r.size, 10), np.float32)
    pdf = np.zeros(r.size, np.float32)
    check(lib().er_test_hdri_sample(rm.ctx, ptr(r), C.c_size_t(r.size), ptr(out), ptr(pdf)))
    want, wpdf = rs.hdri_sample(r)
    # texel choice, inverse-transformed uv and texel value: bit-exact (search + integer/float div only)
    assert np.array_equal(out[:, [0, 1, 2, 3, 7, 8, 9]].view(np.uint32), want[:, [0, 1, 2, 3, 7, 8, 9]].view(np.uint32))
    # direction (sin/cos) and pdf (sin): few ulp
    assert np.abs(out[:, 4:7] - want[:, 4:7]).max() < 2e-6
    fin = np.isfinite(wpdf)
    assert np.array_equal(np.isfinite(pdf), fin)
    assert common.rel_err(pdf[fin], wpdf[fin]).max() < 1e-5


def _random_hitdata(rng, n):
    hd = np.zeros((n, 37), np.float32)
    hd[:, 0] = rng.random(n)            # metallic
    hd[:, 1] = rng.uniform(0.02, 1, n)  # roughness
    hd[:, 2] = rng.random(n)            # clearcoatGloss
    hd[:, 3] = rng.random(n) * 2        # clearcoat
    hd[:, 4] = rng.random(n)            # anisotropic
    hd[:, 5] = 1.5
    hd[:, 6] = np.where(rng.random(n) < 0.1, 1.0, rng.random(n) * 0.5)  # transmission
    hd[:, 7] = rng.random(n)            # specular
    hd[:, 8] = rng.random(n)
    hd[:, 9] = rng.random(n)
    hd[:, 10] = rng.random(n)           # subsurface
    hd[:, 11] = rng.random(n)           # sheen
    hd[:, 12] = 1.0
    hd[:, 15:18] = rng.random((n, 3))   # emission
    hd[:, 18:21] = rng.random((n, 3))   # albedo
    hd[:8, 18:21] = 0.0                 # Cdlum == 0 branch
    nrm = rng.normal(size=(n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    a = np.where(np.abs(nrm[:, 0:1]) < 0.9, [[1.0, 0, 0]], [[0, 1.0, 0]])
    t = np.cross(a, nrm)
    t /= np.linalg.norm(t, axis=1, keepdims=True)
    b = np.cross(nrm, t)
    hd[:, 25:28] = nrm
    hd[:, 28:31] = nrm
    hd[:, 31:34] = t
    hd[:, 34:37] = b
    return hd, nrm


def test_disney(gpu, ref):
    import ctypes as C
    from elevenrender_b200.capi import check, lib, ptr
    rm = host.RenderingManager()
    rng = np.random.default_rng(11)
    n = 6000
    hd, nrm = _random_hitdata(rng, n)

    def hemi(flip_frac):
        v = rng.normal(size=(n, 3))
        v /= np.linalg.norm(v, axis=1, keepdims=True)
        s = np.sign(np.sum(v * nrm, axis=1, keepdims=True))
        v = v * s
        flip = rng.random(n) < flip_frac
        v[flip] *= -1
        return v.astype(np.float32)

    V, L = hemi(0.05), hemi(0.15)
    r3 = rng.random((n, 3)).astype(np.float32)
    s = np.zeros((n, 3), np.float32)
    e = np.zeros((n, 3), np.float32)
    p = np.zeros(n, np.float32)
    check(lib().er_test_disney(rm.ctx, ptr(hd), ptr(V), ptr(L), ptr(r3), C.c_size_t(n), ptr(s), ptr(e), ptr(p)))
    ws, we, wp = np.zeros_like(s), np.zeros_like(e), np.zeros_like(p)
    for i in range(n):
        ws[i], we[i], wp[i] = ref.disney(hd[i], V[i], L[i], r3[i])
    # every operation is IEEE (no contraction) or the host libm's bits (er_libm.cuh): the same words, except where
    # the restated sinf/cosf differ from glibc's FMA build (~1e-8 of calls)
    fin = np.isfinite(we).all(axis=1) & np.isfinite(wp)
    assert np.array_equal(np.isfinite(e).all(axis=1) & np.isfinite(p), fin)
    same = lambda a, b: (a.view(np.uint32) == b.view(np.uint32)) | ((a == 0) & (b == 0)) | (np.isnan(a) & np.isnan(b))
    for what, a, b in (("sample", s, ws), ("eval", e, we), ("pdf", p, wp)):
        bad = int((~same(a, b)).sum())
        assert bad <= max(1, a.size // 100000), f"DisneySample/Eval/Pdf {what}: {bad} of {a.size} words differ"
    assert np.abs(s - ws).max() < 5e-6
    assert common.rel_err(e[fin], we[fin], floor=1e-6).max() < 2e-4
    assert common.rel_err(p[fin], wp[fin], floor=1e-6).max() < 2e-4
    # the exact-zero structure (back faces, transmission >= 1) is identical
    assert np.array_equal(e == 0, we == 0)
    rm.close()


def test_generate_hitdata(pair):
    import ctypes as C
    from elevenrender_b200.capi import check, lib, ptr, HIT_DTYPE
    spec, rm, rs = pair
    rays = common.primary_rays(spec)
    hits = rm.trace(rays)
    hits = hits[hits["valid"] != 0]
    assert hits.size > 500
    out = np.zeros((hits.size, 37), np.float32)
    check(lib().er_test_hitdata(rm.ctx, ptr(hits), C.c_size_t(hits.size), ptr(out)))
    words = common.hits_struct_to_words(hits).view(np.float32)
    mats_seen = set()
    for i in range(0, hits.size, 3):
        want = rs.generate_hitdata(words[i])
        mats_seen.add(int(hits["material_id"][i]))
        # everything but roughness/metallic (powf) is exact
        exact = [k for k in range(37) if k not in (0, 1)]
        assert np.array_equal(out[i, exact].view(np.uint32), want[exact].view(np.uint32)), f"hit {i}"
        assert common.rel_err(out[i, :2], want[:2], floor=1e-7).max() < 1e-5
    assert len(mats_seen) >= 12


def test_camera_rays(gpu, ref):
    for name in ("grid", "grid_rot"):
        spec = common.make_spec(name)
        rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"]))
        rm.upload(common.to_ours(spec))
        rs = common.to_ref(spec)
        rng = np.random.default_rng(2)
        n = 600
        xy = np.stack([rng.integers(0, spec["x_res"], n), rng.integers(0, spec["y_res"], n)], axis=1).astype(np.int32)
        r5 = rng.random((n, 5)).astype(np.float32)
        got = rm.camera_rays(xy, r5)
        want = np.stack([rs.camera_ray(int(xy[i, 0]), int(xy[i, 1]), r5[i]) for i in range(n)])
        if name == "grid":  # no rotation, no lens: sin(0)/cos(0) exact -> bit-exact rays
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        else:  # rotation sin/cos come from the host libm on both sides; lens sampling uses device sin/cos
            assert np.abs(got - want).max() < 1e-6
        rm.close()
        rs.close()
