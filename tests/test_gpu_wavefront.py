"""-m gpu: the wavefront render kernel (er_wavefront.cuh) under scheduling stress. Whatever the warp-role
thresholds, queue pressure and suspend/resume traffic, every pixel must get exactly the samples of the
one-thread-per-pixel kernel: passes, RNG state and sample counters bit-identical."""
import os

import numpy as np
import pytest

import common
from elevenrender_b200 import capi, host

pytestmark = pytest.mark.gpu


def render(spec, spp, kernel, spp_per_launch, max_depth=5):
    pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR, kernel=kernel,
                                 spp_per_launch=spp_per_launch, max_depth=max_depth)
    rm = host.RenderingManager(pars)
    rm.upload(common.to_ours(spec))
    rm.render_samples(spp)
    passes = np.stack([rm.get_pass(p) for p in ("beauty", "normal", "tangent", "bitangent")])
    rng, sa = rm.get_pixel_state()
    rm.close()
    return passes, rng, sa


def big_spec(name, x_res, y_res):
    s = common.make_spec(name)
    s["x_res"], s["y_res"] = x_res, y_res
    return s


@pytest.fixture(scope="module")
def want(gpu):
    out = {}
    for name in ("grid", "spheres", "textured"):
        spec = big_spec(name, 320, 200)
        out[name] = (spec, render(spec, 6, kernel=1, spp_per_launch=6))
    return out


# (thr_descend, thr_switch): defaults, eager suspension, never-break descent, always-break descent
THRESHOLDS = [(20, 12), (24, 20), (32, 32), (1, 1), (8, 31), (28, 4)]


@pytest.mark.parametrize("thr", THRESHOLDS)
@pytest.mark.parametrize("name", ["grid", "spheres", "textured"])
def test_wavefront_equals_pixel_kernel(want, name, thr):
    spec, (p1, r1, s1) = want[name]
    os.environ["ER_WF_THR_DESCEND"], os.environ["ER_WF_THR_SWITCH"] = str(thr[0]), str(thr[1])
    os.environ["ER_WF_WATCHDOG_MS"] = "5000"
    try:
        for rep in range(2):
            p0, r0, s0 = render(spec, 6, kernel=0, spp_per_launch=4)
            assert np.array_equal(r0, r1), f"{name} {thr}: rng state differs"
            assert np.array_equal(s0, s1), f"{name} {thr}: sample counters differ"
            assert np.array_equal(p0.view(np.uint32), p1.view(np.uint32)), f"{name} {thr}: passes differ"
    finally:
        for k in ("ER_WF_THR_DESCEND", "ER_WF_THR_SWITCH", "ER_WF_WATCHDOG_MS"):
            os.environ.pop(k, None)


def test_wavefront_depth16(gpu):
    spec = big_spec("spheres", 160, 120)
    a = render(spec, 3, kernel=0, spp_per_launch=3, max_depth=16)
    b = render(spec, 3, kernel=1, spp_per_launch=3, max_depth=16)
    for x, y in zip(a, b):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32))


# (thr_batch, coop_min): per-warp batches + every leaf pair tested by the whole warp; deep waves + pairs of
# two or more; the defaults' neighbourhood
BATCH_COOP = [(32, 1), (512, 2), (128, 4), (384, 16)]


@pytest.mark.parametrize("bc", BATCH_COOP)
@pytest.mark.parametrize("name", ["spheres", "textured"])
def test_batch_threshold_and_cooperative_leaves(want, name, bc):
    """shading waves (ER_WF_THR_BATCH) and warp-cooperative leaf tests (ER_WF_COOP_MIN) change the
    schedule, never a result"""
    spec, (p1, r1, s1) = want[name]
    os.environ["ER_WF_THR_BATCH"], os.environ["ER_WF_COOP_MIN"] = str(bc[0]), str(bc[1])
    os.environ["ER_WF_WATCHDOG_MS"] = "5000"
    try:
        p0, r0, s0 = render(spec, 6, kernel=0, spp_per_launch=4)
        assert np.array_equal(r0, r1) and np.array_equal(s0, s1)
        assert np.array_equal(p0.view(np.uint32), p1.view(np.uint32))
    finally:
        for k in ("ER_WF_THR_BATCH", "ER_WF_COOP_MIN", "ER_WF_WATCHDOG_MS"):
            os.environ.pop(k, None)


@pytest.mark.parametrize("coop", [1, 4, 1000])
def test_dense_leaves(gpu, coop):
    """1.2 M triangles (several per leaf, candidates that need the exact-distance comparison): the
    wavefront kernel with warp-cooperative, mixed and per-lane leaf tests equals the per-pixel kernel"""
    spec = big_spec("dense", 160, 100)
    os.environ["ER_WF_COOP_MIN"] = str(coop)
    os.environ["ER_WF_WATCHDOG_MS"] = "20000"
    try:
        a = render(spec, 3, kernel=0, spp_per_launch=3)
    finally:
        os.environ.pop("ER_WF_COOP_MIN", None)
        os.environ.pop("ER_WF_WATCHDOG_MS", None)
    b = render(spec, 3, kernel=1, spp_per_launch=3)
    for x, y in zip(a, b):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32))


@pytest.mark.parametrize("chunk", [1, 2, 5])
@pytest.mark.parametrize("name", ["grid", "spheres", "textured"])
def test_sample_chunks(want, name, chunk):
    """a pixel's samples are handed out in chunks (ER_WF_CHUNK_SPP) that different slots and SMs render one
    after the other, the RNG stream and running means carried through global memory: same bits"""
    spec, (p1, r1, s1) = want[name]
    os.environ["ER_WF_CHUNK_SPP"] = str(chunk)
    os.environ["ER_WF_CHUNKED"] = "1"
    os.environ["ER_WF_WATCHDOG_MS"] = "5000"
    try:
        p0, r0, s0 = render(spec, 6, kernel=0, spp_per_launch=6)
        assert np.array_equal(r0, r1) and np.array_equal(s0, s1)
        assert np.array_equal(p0.view(np.uint32), p1.view(np.uint32))
    finally:
        os.environ.pop("ER_WF_CHUNK_SPP", None)
        os.environ.pop("ER_WF_CHUNKED", None)
        os.environ.pop("ER_WF_WATCHDOG_MS", None)


def test_sample_chunks_more_slots_than_pixels(gpu):
    """1536 pixels, 2048 slots, one sample per chunk: most tickets are drawn while the pixel's previous
    chunk is still in flight, so slots wait for the hand-over all the time"""
    spec = common.make_spec("spheres")
    os.environ["ER_WF_CHUNK_SPP"] = "1"
    os.environ["ER_WF_CHUNKED"] = "1"
    os.environ["ER_WF_WATCHDOG_MS"] = "10000"
    try:
        a = render(spec, 12, kernel=0, spp_per_launch=12)
    finally:
        os.environ.pop("ER_WF_CHUNK_SPP", None)
        os.environ.pop("ER_WF_CHUNKED", None)
        os.environ.pop("ER_WF_WATCHDOG_MS", None)
    b = render(spec, 12, kernel=1, spp_per_launch=12)
    for x, y in zip(a, b):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32))


def test_bounds_check_build(gpu):
    """the -DWF_CHECK twin of the library (every slot, queue, triangle, pixel and record index the
    scheduler hands around is bounds-checked; a violation fails the launch) renders through every
    scheduling path without a violation"""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = os.path.join(root, "elevenrender_b200", "liber_cuda_check.so")
    assert os.path.exists(lib), "liber_cuda_check.so missing: run __graft_entry__.build()"
    env = dict(os.environ, ER_CUDA_LIB=lib)
    r = subprocess.run([sys.executable, os.path.join(root, "profiles", "sanitize_run.py")], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0 and r.stdout.strip().endswith("done"), r.stdout[-2000:] + r.stderr[-2000:]


def test_block_order_does_not_change_the_image(gpu):
    """er_set_block_order: the row blocks handed out in another order (reversed, shuffled) give the same frame bit
    for bit; a wrong-sized or non-permutation order is refused / ignored"""
    spec = common.make_spec("spheres")
    spec["x_res"], spec["y_res"] = 160, 100  # 13 row blocks of 8, the last one partial
    spp = 4
    frames = []
    for order in (None, list(range(12, -1, -1)), list(np.random.default_rng(3).permutation(13))):
        rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR))
        rm.upload(common.to_ours(spec))
        rm.set_block_order(order)
        rm.render_samples(spp)
        frames.append(rm.get_pass("beauty"))
        with pytest.raises(capi.ErError):
            rm.set_block_order([0, 0, 1])
        rm.set_block_order([1, 0])  # wrong size for this shard: ignored at launch
        rm.reset()
        rm.render_samples(spp)
        assert np.array_equal(rm.get_pass("beauty").view(np.uint32), frames[-1].view(np.uint32))
        rm.close()
    assert np.array_equal(frames[0].view(np.uint32), frames[1].view(np.uint32))
    assert np.array_equal(frames[0].view(np.uint32), frames[2].view(np.uint32))
    # on a stripe: blocks are counted from the stripe's first block
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR))
    rm.upload(common.to_ours(spec))
    rm.set_stripe(24, 72, 8)
    rm.set_block_order([5, 3, 1, 0, 2, 4])
    rm.render_samples(spp)
    got = rm.get_pass("beauty")
    assert np.array_equal(got[24:72].view(np.uint32), frames[0][24:72].view(np.uint32))
    assert np.all(got[:24] == 0) and np.all(got[72:] == 0)
    rm.close()
