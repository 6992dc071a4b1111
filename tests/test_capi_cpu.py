"""CPU tests: the C-ABI library loads, exports every symbol include/er_cuda.h declares, and fails
loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import common
from elevenrender_b200 import capi

HEADER = os.path.join(common.ROOT, "include", "er_cuda.h")


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"ER_API\s+[\w\s\*]+?\b(er_\w+)\s*\(", src)))


def test_header_symbols_exported():
    names = declared_symbols()
    assert len(names) >= 35
    L = C.CDLL(capi.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"{n} declared in er_cuda.h but not exported"
    assert sorted(capi.EXPORTS) == names, "capi.EXPORTS and include/er_cuda.h disagree"


def test_record_sizes():
    assert capi.NODE_DTYPE.itemsize == 44 and capi.HIT_DTYPE.itemsize == 84
    assert C.sizeof(capi.ErRenderOpts) == 32 and C.sizeof(capi.ErCounters) == 56


def test_no_cpu_fallback():
    """without a GPU every compute entry point must fail with ER_ERR_NO_DEVICE, never compute"""
    try:
        n = capi.device_count()
    except capi.ErError as e:
        assert e.code == -2
        n = 0
    if n > 0:
        pytest.skip("a CUDA device is present")
    ctx = C.c_void_p()
    rc = capi.lib().er_create(0, C.byref(ctx))
    assert rc == -2 and not ctx.value
    assert b"CUDA" in capi.lib().er_last_error()
    from elevenrender_b200 import host
    with pytest.raises(capi.ErError):
        host.RenderingManager()


def test_invalid_arguments():
    L = capi.lib()
    assert L.er_bvh_build(None, C.c_size_t(3), C.c_size_t(156), None, None, 1) == -1
    assert L.er_hdri_build_cdf(None, None, None) == -1
    assert L.er_render_poll(None, None, None) == -1
    t = np.zeros((2, 39), dtype=np.float32)
    nodes = np.zeros(capi.ER_BVH_NODE_COUNT, dtype=capi.NODE_DTYPE)
    idx = np.zeros(2, dtype=np.int32)
    assert L.er_bvh_build(capi.ptr(t), C.c_size_t(2), C.c_size_t(8), capi.ptr(nodes), capi.ptr(idx), 1) == -1


def test_empty_scene_bvh():
    nodes, idx = capi.bvh_build(np.zeros((0, 39), dtype=np.float32), 1)
    assert idx.size == 0
    assert np.all(nodes["from"] == 0) and np.all(nodes["to"] == 0)
    assert np.all(nodes["idx"] == np.arange(capi.ER_BVH_NODE_COUNT))
