"""N4 (-m gpu): the off-by-default extensions of the estimator (ErRenderOpts.extensions). Each changes the image
relative to the reference, so each is checked against a CPU restatement of what it must compute:

  ER_EXT_MIS           E[default] = 2 E[MIS] on primary-hit pixels of an environment-lit scene: the reference adds
                       the environment once through NEE and once more through the BRDF-sampled ray (kernel.cpp:
                       517-522,567-575, the weights of :571-572 are dropped); weighted, it is counted once
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


def test_mis_counts_the_environment_once(gpu):
    spec = common.make_spec("grid")
    spec["hdri"] = (32, 16, 3, soft_env(), False)
    spec["x_res"], spec["y_res"] = 96, 64
    spp = 192
    base, n = render(spec, spp)
    mis, _ = render(spec, spp, ext=host.ER_EXT_MIS)
    hit = (n != 0).any(axis=-1)
    assert hit.mean() > 0.3
    ratio = base[hit].mean() / mis[hit].mean()
    # 2 up to the last bounce (its NEE term has no BRDF-sampled partner) and texel row 0 (infinite pdf: NEE adds 0)
    assert 1.85 < ratio < 2.05, ratio
    miss = ~hit
    if miss.any():  # camera rays that leave the scene see the environment once in both
        assert abs(base[miss].mean() / mis[miss].mean() - 1.0) < 0.02


def wall_scene(light_radiance=(6.0, 5.0, 4.0), occluder=False):
    s = dict(materials=[], textures=[], hdri=(0.0, 0.0, 0.0), x_res=48, y_res=32)
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
    assert np.all(base == 0)  # black environment, no emission: without the extension the image is black
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
        assert np.allclose(got, want, rtol=0.02), (x, y, got, want)
    rs.close()
    # linear in the radiance: doubling it doubles every sample (an exact scaling in binary)
    img2, _ = render(wall_scene(light_radiance=(12.0, 10.0, 8.0)), spp, ext=host.ER_EXT_POINT_LIGHTS)
    assert np.allclose(img2, 2.0 * img, rtol=1e-6)
    # an occluder between the light and the wall: its shadow is black, the rest unchanged
    occ, n_occ = render(wall_scene(occluder=True), spp, ext=host.ER_EXT_POINT_LIGHTS)
    assert occ[16, 24].max() == 0.0 or occ[16, 24].max() < img[16, 24].max() * 0.05


def test_roulette_is_unbiased(gpu):
    spec = common.make_spec("cornell")
    spec["x_res"], spec["y_res"] = 64, 64
    spp = 256
    a, _ = render(spec, spp)
    b, _ = render(spec, spp, ext=host.ER_EXT_ROULETTE)
    assert not np.array_equal(a, b)
    assert abs(b.mean() / a.mean() - 1.0) < 0.015, (a.mean(), b.mean())
