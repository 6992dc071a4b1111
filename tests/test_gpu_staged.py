"""The staged render path (kernel 3: k_st_logic / k_st_trace / k_st_shade waves, er_staged.cuh) against the
persistent kernel (kernel 0) and the oracle (-m gpu). Every pixel runs the same operations in the same
order in both, so passes, RNG states and sample counters must be identical bit for bit."""
import os

import numpy as np
import pytest

import common
from elevenrender_b200 import host

pytestmark = pytest.mark.gpu
SCENES = ["grid", "grid_rot", "cornell", "textured", "spheres", "miss", "dense"]


def render(spec, spp, kernel, shard=(0, 1, 8), spp_per_launch=0, max_depth=5, order=host.ER_RNG_RTL, stripe=None):
    pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=order, kernel=kernel,
                                 spp_per_launch=spp_per_launch, max_depth=max_depth)
    rm = host.RenderingManager(pars, shard=shard)
    rm.upload(common.to_ours(spec))
    if stripe is not None:
        rm.set_stripe(stripe[0], stripe[1], 8)
    rm.render_samples(spp)
    passes = np.stack([rm.get_pass(p) for p in ("beauty", "denoise", "normal", "tangent", "bitangent")])
    rng, sa = rm.get_pixel_state()
    rm.close()
    return passes, rng, sa


def same(a, b):
    return all(np.array_equal(x.view(np.uint32), y.view(np.uint32)) for x, y in zip(a, b))


@pytest.mark.parametrize("name", SCENES)
def test_staged_equals_persistent(gpu, name):
    spec = common.make_spec(name)
    for spp, order in ((1, host.ER_RNG_RTL), (5, host.ER_RNG_RTL), (8, host.ER_RNG_LTR)):
        a = render(spec, spp, 0, order=order)
        b = render(spec, spp, 3, order=order)
        assert same(a, b), f"{name} spp={spp}: staged path differs from the persistent kernel"


def test_staged_matches_golden(gpu):
    for name in ("grid", "textured"):
        g = np.load(os.path.join(common.GOLDEN_DIR, f"{name}.npz"))
        spec = common.make_spec(name)
        passes, rng, sa = render(spec, 4, 3)
        err = common.rel_err(passes[[0, 2, 3, 4], ..., :3], g["passes_4"][..., :3], floor=1e-5).max(axis=-1)
        assert (err > 1e-3).mean() <= 1e-4
        assert np.array_equal(sa, g["samples_4"])


def test_staged_launch_splits_and_depth(gpu):
    spec = common.make_spec("spheres")
    whole = render(spec, 12, 3)
    split = render(spec, 12, 3, spp_per_launch=5)  # 5 + 5 + 2: the pixel records carry the state across launches
    assert same(whole, split)
    deep0 = render(spec, 3, 0, max_depth=16)
    deep3 = render(spec, 3, 3, max_depth=16)
    assert same(deep0, deep3)


def test_staged_shards(gpu):
    spec = common.make_spec("grid")
    full = render(spec, 4, 3)
    acc = np.zeros_like(full[0])
    for k in range(3):
        part = render(spec, 4, 3, shard=(k, 3, 8))
        acc += part[0]
    assert np.array_equal(acc[[0, 2, 3, 4]].view(np.uint32), full[0][[0, 2, 3, 4]].view(np.uint32))
    h = spec["y_res"]
    cut = (h // 2) // 8 * 8
    top = render(spec, 4, 3, stripe=(0, cut))
    bot = render(spec, 4, 3, stripe=(cut, h))
    both = top[0] + bot[0]
    assert np.array_equal(both[[0, 2, 3, 4]].view(np.uint32), full[0][[0, 2, 3, 4]].view(np.uint32))
