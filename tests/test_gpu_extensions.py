"""N4 (-m gpu): the off-by-default extensions of the estimator (ErRenderOpts.extensions). Each changes the image
relative to the reference, so each is checked against a CPU restatement of what it must compute:

  ER_EXT_MIS           one-bounce paths at 1 spp restated per pixel from the oracle's own functions and the pixel's RNG
                       stream: NEE weighted hw, the environment seen by the BRDF-sampled ray weighted bw
                       (kernel.cpp:571-572, computed and dropped there)
  ER_EXT_POINT_LIGHTS  one diffuse wall, a black environment, one light: the pixel value is
                       radiance / d^2 * DisneyEval * |cos| * 2 pi (pointLight(), kernel.cpp:269-301), with
                       DisneyEval and the hit taken from the ORACLE (BVH::transverse, generateHitData, DisneyEval);
                       linear in the radiance (bit-exact doubling); an occluder removes it
  ER_EXT_ROULETTE      unbiased: the mean radiance of the Cornell box is unchanged
and with extensions = 0 nothing changes (bit-identical to the default render)."""
import numpy as np
import pytest

import common
from elevenrender_b200 import host, scenes

pytestmark = pytest.mark.gpu


def render(spec, spp, ext=0, order=host.ER_RNG_LTR, kernel=0):
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=order, extensions=ext,
                                                     kernel=kernel, spp_per_launch=spp))
    rm.upload(common.to_ours(spec))
    rm.render_samples(spp)
    b, n = rm.get_pass("beauty"), rm.get_pass("normal")
    rm.close()
    return b[..., :3].astype(np.float64) * (spp + 1) / spp, n[..., :3]  # running mean starts its counter at 1 (F3)


def soft_env(w=32, h=16):
    y, x = np.mgrid[0:h, 0:w]
    v = 0.45 + 0.15 * np.sin(2 * np.pi * x / w) * np.cos(np.pi * y / h)
    return np.repeat(v[..., None], 3, axis=2).astype(np.float32).reshape(-1)


def test_extensions_off_is_the_default(gpu):
    spec = common.make_spec("spheres")
    a, _ = render(spec, 3)
    b, _ = render(spec, 3, ext=0, kernel=1)
    assert np.array_equal(a, b)


def test_mis_weights_match_cpu_restatement(gpu, ref):
    """One sample of one-bounce paths restated on the CPU from the oracle's own functions (camera ray, BVH::transverse,
    generateHitData, HDRI::sample/pdf, DisneySample/Eval/Pdf, environment lookup) with the pixel's RNG stream: a wall in
    front of an environment map, so every path is hit -> (NEE, BRDF-sampled ray that leaves the scene). The default
    estimator adds NEE + environment; ER_EXT_MIS weights them hw / bw (kernel.cpp:571-572). Both images are checked."""
    spec = wall_scene()
    spec["hdri"] = (32, 16, 3, soft_env(), False)
    spec["point_lights"] = []
    w, h = spec["x_res"], spec["y_res"]
    base, _ = render(spec, 1)
    mis, _ = render(spec, 1, ext=host.ER_EXT_MIS)
    base, mis = base / 2.0, mis / 2.0  # undo render()'s (spp+1)/spp: compare the stored running mean light/(1+1)
    rs = common.to_ref(spec)
    rng = np.random.default_rng(3)
    checked = 0
    for _ in range(60):
        x, y = int(rng.integers(0, w)), int(rng.integers(0, h))
        _, u = ref.rng_stream(y * w + x, 10)
        ray = rs.camera_ray(x, y, u[0:5])  # LTR: the five camera draws in order
        hit = rs.trace(ray.reshape(1, 6))[0]
        if not hit[15:16].view(np.int32)[0]:
            continue
        hd = rs.generate_hitdata(hit)
        assert u[5] <= hd[12]  # opaque
        pos, nrm = hd[22:25], hd[25:28]
        wo = -ray[3:6]
        o10, hpdf = rs.hdri_sample(u[6:7])
        wih, val, hdripdf = o10[0, 4:7], o10[0, 7:10], float(hpdf[0])
        wib, evalh, _ = ref.disney(hd, wo, wih, u[7:10])
        _, evalb, brdfpdf = ref.disney(hd, wo, wib, u[7:10])
        sh = rs.trace(np.concatenate([pos + nrm * np.float32(0.001), wih]).astype(np.float32).reshape(1, 6))[0]
        blocked = bool(sh[15:16].view(np.int32)[0]) and sh[20:21].view(np.int32)[0] != hit[20:21].view(np.int32)[0]
        with np.errstate(all="ignore"):
            hdri_int = (np.zeros(3) if blocked else val.astype(np.float64)) * evalh * abs(float(np.dot(wih, nrm))) / hdripdf
            factor = evalb.astype(np.float64) * abs(float(np.dot(wib, nrm))) / brdfpdf
        nd = wib / np.linalg.norm(wib)
        nxt = rs.trace(np.concatenate([hit[0:3] + wib * np.float32(0.001), nd]).astype(np.float32).reshape(1, 6))[0]
        if nxt[15:16].view(np.int32)[0]:
            continue  # (cannot happen on a single wall)
        uv = ref.spherical_mapping((-nd).reshape(1, 3))
        env = rs.texture_lookup(-1, 3, uv)[0].astype(np.float64)
        hw = hdripdf / (brdfpdf + hdripdf)
        bw = brdfpdf / (brdfpdf + hdripdf)
        if not (0.0 <= hw <= 1.0):
            hw, bw = 0.0, 1.0
        want_base = np.clip(hdri_int + factor * env, 0, 10) / 2.0
        want_mis = np.clip(hdri_int * hw + factor * bw * env, 0, 10) / 2.0
        if not (np.isfinite(want_base).all() and np.isfinite(want_mis).all()):
            continue
        assert np.allclose(base[y, x], want_base, rtol=2e-4, atol=1e-6), (x, y, base[y, x], want_base)
        assert np.allclose(mis[y, x], want_mis, rtol=2e-4, atol=1e-6), (x, y, mis[y, x], want_mis, hw, bw)
        checked += 1
    rs.close()
    assert checked >= 30


def wall_scene(light_radiance=(6.0, 5.0, 4.0), occluder=False):
    # a 1x1 environment of 1e-3: its NEE pdf is infinite (HDRI.cpp:100-107, sin(0)), so NEE adds exactly 0 and a miss adds
    # reduction * 1e-3 (a black one would make every sample NaN: 0/0 in HDRI::pdf, and NaN samples are discarded)
    s = dict(materials=[], textures=[], hdri=(1e-3, 1e-3, 1e-3), x_res=48, y_res=32)
    z = 3.0
    v = [[[-4, -3, z], [4, -3, z], [4, 3, z]], [[-4, -3, z], [4, 3, z], [-4, 3, z]]]
    if occluder:
        v += [[[-0.5, -0.5, 2.0], [0.5, -0.5, 2.0], [0.5, 0.5, 2.0]], [[-0.5, -0.5, 2.0], [0.5, 0.5, 2.0], [-0.5, 0.5, 2.0]]]
    v = np.float32(v)
    s["tris"] = scenes.pack_tris(v, np.broadcast_to(np.float32([0, 0, -1]), v.shape))
    s["camera"] = dict(position=(0.0, 0.0, -1.0))
    s["point_lights"] = [((0.0, 0.0, 1.0), tuple(light_radiance))]  # on the axis, behind the occluder as seen from the wall
    return s


def test_point_light_matches_cpu_restatement(gpu, ref):
    spec = wall_scene()
    spp = 64
    img, n = render(spec, spp, ext=host.ER_EXT_POINT_LIGHTS)
    base, _ = render(spec, spp)
    assert base.max() < 2e-3  # without the extension only the 1e-3 environment is seen
    # CPU restatement for pixel centres: the oracle's hit, hit data and DisneyEval; the light term by hand
    rs = common.to_ref(spec)
    px = [(24, 16), (10, 8), (40, 25), (5, 30)]
    lp, lr = np.float32(spec["point_lights"][0][0]), np.float32(spec["point_lights"][0][1])
    for (x, y) in px:
        want = np.zeros(3)
        m = 6  # average the restatement over the pixel footprint like the jittered samples do
        for jy in range(m):
            for jx in range(m):
                r5 = np.float32([(jx + 0.5) / m, (jy + 0.5) / m, 0.5, 0.5, 0.5])  # r1, r2 jitter the sensor point (kernel.cpp:384-385)
                ray = rs.camera_ray(x, y, r5)
                hitw = rs.trace(ray.reshape(1, 6))[0]
                hd = rs.generate_hitdata(hitw)
                pos, nrm = hitw[0:3], hitw[3:6]
                wi = (lp - pos) / np.linalg.norm(lp - pos)
                d = np.linalg.norm(lp - pos)
                wo = -ray[3:6] / np.linalg.norm(ray[3:6])
                _, ev, _ = ref.disney(hd, wo, wi, np.float32([0.1, 0.2, 0.3]))
                want += lr / (d * d) * ev * abs(float(np.dot(wi, nrm))) * (2.0 * np.pi)
        want /= m * m
        got = img[y, x]
        assert np.allclose(got, want, rtol=0.02, atol=2e-3), (x, y, got, want)
    rs.close()
    # linear in the radiance: doubling it doubles the light term of every sample (the same paths: the same RNG
    # stream); what does not double is the 1e-3 environment seen by the bounce rays (~5e-4 per pixel)
    img2, _ = render(wall_scene(light_radiance=(12.0, 10.0, 8.0)), spp, ext=host.ER_EXT_POINT_LIGHTS)
    assert np.allclose(img2, 2.0 * img, rtol=0, atol=1.5e-3) and img.min() > 0.05
    # an occluder between the light and the wall: its shadow is black, the rest unchanged
    occ, n_occ = render(wall_scene(occluder=True), spp, ext=host.ER_EXT_POINT_LIGHTS)
    # the light (z = 1) throws the occluder's shadow (z = 2, half-width 0.5) onto the wall (z = 3) with half-width 1; the
    # camera (z = -1) sees the wall beyond |x| = 0.67 past the occluder's edge: pixel column 34 looks at x ~ 0.85, in shadow
    assert img[16, 34].max() > 0.3 and occ[16, 34].max() < 0.02 * img[16, 34].max()
    assert np.allclose(occ[16, 46], img[16, 46], rtol=1e-3)  # outside the shadow: unchanged


def test_roulette_is_unbiased(gpu):
    spec = common.make_spec("cornell")
    spec["x_res"], spec["y_res"] = 64, 64
    spp = 256
    a, _ = render(spec, spp)
    b, _ = render(spec, spp, ext=host.ER_EXT_ROULETTE)
    assert not np.array_equal(a, b)
    assert abs(b.mean() / a.mean() - 1.0) < 0.015, (a.mean(), b.mean())
