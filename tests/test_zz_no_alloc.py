"""(-m gpu) er_trace_noalloc — the nearest-hit query through the library's fixed device staging, for callers that
cannot call cudaMalloc. No counterpart in the reference: the check is
against the existing entry point, er_trace, on the same rays — the whole 84-byte Hit record bit-identical, the
tolerance tests/test_gpu_trace.py uses — on two calls in a row (the staging keeps the last call's contents: the
second, SMALLER batch must not see the first one's hits), then the size limit.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

import common
from elevenrender_b200 import capi, host, scenes

pytestmark = pytest.mark.gpu


def _words(h):
    return common.hits_struct_to_words(h)


def test_trace_noalloc_matches_trace_twice(gpu):
    spec = common.make_spec("cornell")
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"]))
    rm.upload(common.to_ours(spec))
    lo = spec["tris"][:, 0:9].reshape(-1, 3).min(axis=0) - 0.5
    hi = spec["tris"][:, 0:9].reshape(-1, 3).max(axis=0) + 0.5
    first = scenes.random_rays(30000, origin_box=(lo, hi), target_box=(lo, hi), seed=5)
    second = scenes.random_rays(9000, origin_box=(lo, hi), target_box=(lo + 2.0, hi + 2.0), seed=6)  # many misses
    for rays in (first, second):
        want = _words(rm.trace(rays))
        got = _words(rm.trace_noalloc(rays))
        bad = np.nonzero((got != want).any(axis=1))[0]
        assert bad.size == 0, f"{bad.size} of {got.shape[0]} hits differ from er_trace; first {bad[:5]}"
    assert (rm.trace(first)["valid"] != 0).sum() > 1000  # the comparison saw real hits
    # the primary rays of the frame too (coherent rays, golden vectors of the oracle)
    rays = common.primary_rays(spec)[: capi.TRACE_NOALLOC_MAX]
    assert (_words(rm.trace_noalloc(rays)) == _words(rm.trace(rays))).all()
    rm.close()


def test_trace_noalloc_limit_and_empty(gpu):
    spec = common.make_spec("grid")
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"]))
    rm.upload(common.to_ours(spec))
    assert rm.trace_noalloc(np.zeros((0, 6), np.float32)).shape == (0,)
    full = scenes.random_rays(capi.TRACE_NOALLOC_MAX, seed=9)
    assert (_words(rm.trace_noalloc(full)) == _words(rm.trace(full))).all()  # exactly the maximum
    with pytest.raises(capi.ErError) as e:
        rm.trace_noalloc(np.zeros((capi.TRACE_NOALLOC_MAX + 1, 6), np.float32))
    assert "ER_TRACE_NOALLOC_MAX" in str(e.value) and str(capi.TRACE_NOALLOC_MAX) in str(e.value)
    # the context still works after the refusal
    assert (_words(rm.trace_noalloc(full[:100])) == _words(rm.trace(full[:100]))).all()
    rm.close()
