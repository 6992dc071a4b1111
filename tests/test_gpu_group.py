"""Device groups in the C ABI (er_group_*, SURVEY §8b/e: "multi-GPU reduce inside er_get_pass") (-m gpu).
A group shards the frame over its member contexts (scene replicated), renders the shards concurrently and
er_group_get_pass gathers every member's OWN rows straight into the caller's frame. The frame must equal the
single-context frame bit for bit (T6). With one GPU on the box the members are contexts on the same device
(the code path is the same); with two or more they are different devices."""
import numpy as np
import pytest

import common
from elevenrender_b200 import capi, host

pytestmark = pytest.mark.gpu


def single(spec, spp, order):
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=order))
    rm.upload(common.to_ours(spec))
    rm.render_samples(spp)
    out = {p: rm.get_pass(p) for p in ("beauty", "normal", "tangent", "bitangent", "denoise")}
    info = rm.get_render_info()
    rm.close()
    return out, info


@pytest.mark.parametrize("members", [2, 3])
@pytest.mark.parametrize("balanced", [False, True])
def test_group_frame_equals_single_context(gpu, members, balanced):
    ndev = capi.device_count()
    devices = [i % ndev for i in range(members)]
    spec = common.make_spec("spheres")
    spec["x_res"], spec["y_res"] = 96, 72
    spp = 5
    want, info1 = single(spec, spp, host.ER_RNG_LTR)
    g = host.RenderingGroup(host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR), devices)
    g.upload(common.to_ours(spec))
    if balanced:
        yb = g.balance()
        assert yb[0] == 0 and yb[-1] == spec["y_res"] and np.all(np.diff(yb) >= 0)
    g.render_samples(spp)
    assert g.get_render_info() == info1  # min over the members: completed samples + 1
    for p, w in want.items():
        got = g.get_pass(p)
        assert np.array_equal(got.view(np.uint32), w.view(np.uint32)), f"{p}: the group frame differs from the single-context frame"
    g.close()


def test_group_async_start_poll(gpu):
    ndev = capi.device_count()
    spec = common.make_spec("grid")
    spp = 6
    want, _ = single(spec, spp, host.ER_RNG_LTR)
    g = host.RenderingGroup(host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR, spp_per_launch=2),
                            [0, 1 % ndev])
    g.start_rendering(common.to_ours(spec))
    g.wait()
    s, running = g.poll()
    assert s == spp + 1 and not running  # N+1 semantics of dev_samples, as er_render_poll
    assert g.get_render_info()["samples"] == spp + 1
    assert np.array_equal(g.get_pass("beauty").view(np.uint32), want["beauty"].view(np.uint32))
    g.close()


def test_group_rebalance_feedback(gpu):
    """er_group_rebalance: new stripes from the measured kernel times; the frame rendered on them is still the
    single-context frame bit for bit; refused before er_group_balance and before anything was rendered"""
    ndev = capi.device_count()
    devices = [i % ndev for i in range(3)]
    spec = common.make_spec("spheres")
    spec["x_res"], spec["y_res"] = 96, 72
    spp = 5
    want, _ = single(spec, spp, host.ER_RNG_LTR)
    g = host.RenderingGroup(host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR), devices)
    g.upload(common.to_ours(spec))
    with pytest.raises(capi.ErError):
        g.rebalance()  # interleaved blocks: nothing to move
    yb0 = g.balance()
    g.render_samples(spp)
    for _ in range(2):
        yb = g.rebalance()
        assert yb[0] == 0 and yb[-1] == spec["y_res"] and np.all(np.diff(yb) >= 8) and all(int(v) % 8 == 0 for v in yb)
        g.render_samples(spp)  # accumulation restarted by the rebalance
        got = g.get_pass("beauty")
        assert np.array_equal(got.view(np.uint32), want["beauty"].view(np.uint32)), (yb0, yb)
    g.close()
