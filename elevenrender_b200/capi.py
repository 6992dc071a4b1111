"""ctypes binding of the C ABI in include/er_cuda.h (elevenrender_b200/liber_cuda.so).

Thin and literal: every function here is one `extern "C"` entry point.  There is no CPU path behind
it — if the library is missing, or no B200 is present, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# ER_CUDA_LIB selects another build of the same library (A/B runs of kernel variants, WF_CHECK builds)
LIB_PATH = os.environ.get("ER_CUDA_LIB") or os.path.join(_HERE, "liber_cuda.so")

ER_BVH_DEPTH = 18
ER_BVH_NODE_COUNT = (2 << ER_BVH_DEPTH) - 1
ER_PASSES = {"beauty": 0, "denoise": 1, "normal": 2, "tangent": 3, "bitangent": 4}
ER_RNG_LTR, ER_RNG_RTL = 0, 1
ER_EXT_MIS, ER_EXT_POINT_LIGHTS, ER_EXT_ROULETTE = 1, 2, 4
TRACE_NOALLOC_MAX = 65536  # ER_TRACE_NOALLOC_MAX: rays per er_trace_noalloc call

TRI_WORDS = 39   # ErTri: vertices9 uv9 normals9 tangents9 sign objectID materialID
MAT_WORDS = 29   # ErMaterial
CAM_WORDS = 12   # ErCamera
NODE_WORDS = 11  # ErNode
HIT_WORDS = 21   # ErHit

# numpy views of the records (field order of include/er_cuda.h)
NODE_DTYPE = np.dtype([("b1", "<f4", 3), ("b2", "<f4", 3), ("from", "<i4"), ("to", "<i4"), ("idx", "<i4"),
                       ("depth", "<i4"), ("valid", "<i4")])
HIT_DTYPE = np.dtype([("position", "<f4", 3), ("normal", "<f4", 3), ("tangent", "<f4", 3), ("bitangent", "<f4", 3),
                      ("gnormal", "<f4", 3), ("valid", "<i4"), ("objectID", "<u4"), ("material_id", "<u4"),
                      ("tu", "<f4"), ("tv", "<f4"), ("triIdx", "<i4")])
assert NODE_DTYPE.itemsize == 44 and HIT_DTYPE.itemsize == 84


class ErTexture(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("channels", C.c_int32), ("filter", C.c_int32),
                ("data", C.c_void_p)]


class ErSceneDesc(C.Structure):
    _fields_ = [("x_res", C.c_uint32), ("y_res", C.c_uint32), ("tris", C.c_void_p), ("tri_count", C.c_size_t),
                ("tri_stride", C.c_size_t), ("nodes", C.c_void_p), ("node_stride", C.c_size_t),
                ("tri_indices", C.c_void_p), ("materials", C.c_void_p), ("material_count", C.c_int32),
                ("textures", C.POINTER(ErTexture)), ("texture_count", C.c_int32), ("hdri", ErTexture),
                ("hdri_cdf", C.c_void_p), ("hdri_radiance_sum", C.c_float), ("camera", C.c_void_p),
                ("point_lights", C.c_void_p), ("point_light_count", C.c_int32)]


class ErRenderOpts(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("rng_order", C.c_int32), ("spp_per_launch", C.c_int32),
                ("kernel", C.c_int32), ("count_stats", C.c_int32), ("extensions", C.c_int32), ("reserved", C.c_int32 * 2)]


class ErCounters(C.Structure):
    _fields_ = [(n, C.c_ulonglong) for n in
                ("rays", "box_tests", "tri_tests", "candidates", "shaded_hits", "nee_samples", "samples")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class ErDeviceInfo(C.Structure):
    _fields_ = [("name", C.c_char * 256), ("platform", C.c_char * 64), ("memory_bytes", C.c_ulonglong),
                ("max_compute_units", C.c_int32), ("cc_major", C.c_int32), ("cc_minor", C.c_int32),
                ("is_compatible", C.c_int32)]


# every symbol include/er_cuda.h declares
EXPORTS = [
    "er_last_error", "er_device_count", "er_device_info", "er_create", "er_destroy", "er_set_shard", "er_set_stream",
    "er_bvh_build", "er_load_obj", "er_hdri_build_cdf", "er_flatten_roundtrip", "er_upload_scene", "er_set_camera", "er_render_reset", "er_render_samples",
    "er_render_enqueue", "er_render_start", "er_render_poll", "er_render_wait", "er_render_abort", "er_sync",
    "er_get_render_info", "er_get_pass", "er_get_pass_device", "er_get_pixel_state", "er_get_counters",
    "er_reset_counters", "er_get_kernel_time", "er_trace", "er_trace_mode", "er_trace_device", "er_trace_noalloc", "er_camera_rays",
    "er_test_rng", "er_test_geometry", "er_test_libm", "er_test_texture", "er_test_spherical", "er_test_hdri_sample", "er_test_disney", "er_test_hitdata",
    "er_set_shard_stripe", "er_row_cost", "er_balance_stripes", "er_rebalance_stripes", "er_set_block_order", "er_order_blocks", "er_get_pass_owned", "er_host_alloc", "er_host_free",
    "er_host_register", "er_host_unregister",
    "er_group_create", "er_group_destroy", "er_group_size", "er_group_ctx", "er_group_upload_scene", "er_group_balance", "er_group_rebalance",
    "er_group_render_reset", "er_group_render_samples", "er_group_render_start", "er_group_render_poll",
    "er_group_render_wait", "er_group_render_abort", "er_group_get_render_info", "er_group_get_pass",
]

_lib = None


class ErError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"er_cuda error {code}: {msg}")
        self.code = code


def lib():
    """Load liber_cuda.so (built by __graft_entry__.build() / `make -C elevenrender_b200/csrc`)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ErError(-2, f"{LIB_PATH} is missing: build it with `make -C elevenrender_b200/csrc` "
                              "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.er_last_error.restype = C.c_char_p
        for name in EXPORTS:
            fn = getattr(L, name)
            if name != "er_last_error":
                fn.restype = C.c_int
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise ErError(rc, lib().er_last_error().decode(errors="replace"))


def ptr(a):
    """void* of a C-contiguous numpy array (None -> NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def device_count():
    n = C.c_int(0)
    check(lib().er_device_count(C.byref(n)))
    return n.value


def device_info(i):
    info = ErDeviceInfo()
    check(lib().er_device_info(int(i), C.byref(info)))
    return {"name": info.name.decode(), "platform": info.platform.decode(), "memory": int(info.memory_bytes),
            "max_compute_units": int(info.max_compute_units), "cc": (info.cc_major, info.cc_minor),
            "is_compatible": bool(info.is_compatible)}


def bvh_build(tri_words, nthreads=0):
    """er_bvh_build: (n,39) float32 triangle words -> (nodes structured array, tri_indices int32)."""
    tri_words = f32(tri_words).reshape(-1, TRI_WORDS)
    n = tri_words.shape[0]
    nodes = np.zeros(ER_BVH_NODE_COUNT, dtype=NODE_DTYPE)
    idx = np.zeros(max(n, 1), dtype=np.int32)
    check(lib().er_bvh_build(ptr(tri_words), C.c_size_t(n), C.c_size_t(TRI_WORDS * 4), ptr(nodes), ptr(idx),
                             int(nthreads)))
    return nodes, idx[:n]


def hdri_build_cdf(texels, width, height, channels):
    texels = f32(texels)
    t = ErTexture(int(width), int(height), int(channels), 0, texels.ctypes.data)
    cdf = np.zeros(width * height + 1, dtype=np.float32)
    s = C.c_float(0)
    check(lib().er_hdri_build_cdf(C.byref(t), ptr(cdf), C.byref(s)))
    return cdf, float(s.value)


def balance_stripes(cost_per_row, rows_per_block, n):
    """er_balance_stripes: stripe boundaries y[0..n] (multiples of rows_per_block) that split the cost evenly"""
    cost = np.ascontiguousarray(cost_per_row, dtype=np.uint32)
    yb = np.zeros(n + 1, dtype=np.int32)
    check(lib().er_balance_stripes(ptr(cost), int(cost.size), int(rows_per_block), int(n), ptr(yb)))
    return yb


def rebalance_stripes(y_bounds, ms, rows_per_block):
    """er_rebalance_stripes: the feedback step of the balance — new boundaries from the time every stripe took"""
    yb = np.ascontiguousarray(y_bounds, dtype=np.int32)
    t = np.ascontiguousarray(ms, dtype=np.float32)
    assert yb.size == t.size + 1
    out = np.zeros(yb.size, dtype=np.int32)
    check(lib().er_rebalance_stripes(ptr(yb), ptr(t), int(t.size), int(rows_per_block), ptr(out)))
    return out


def order_blocks(cost_per_row, rows_per_block, y_begin, y_end):
    """er_order_blocks: the row blocks of [y_begin, y_end) by descending cost (indices relative to the range)"""
    cost = np.ascontiguousarray(cost_per_row, dtype=np.uint32)
    out = np.zeros(max(1, -(-(y_end - y_begin) // rows_per_block)), dtype=np.int32)
    n = C.c_int(0)
    check(lib().er_order_blocks(ptr(cost), int(cost.size), int(rows_per_block), int(y_begin), int(y_end), ptr(out), C.byref(n)))
    return out[:n.value]


def host_register(array):
    """page-lock a numpy array in place (er_host_register)"""
    check(lib().er_host_register(C.c_void_p(array.ctypes.data), C.c_size_t(array.nbytes)))


def host_unregister(array):
    check(lib().er_host_unregister(C.c_void_p(array.ctypes.data)))


def load_obj(obj_text, mtl_text="", recompute_normals=False):
    """er_load_obj: OBJ + MTL text -> ((n,39) float32 triangle words, [material name per triangle])"""
    ob = obj_text.encode() if isinstance(obj_text, str) else bytes(obj_text)
    mt = mtl_text.encode() if isinstance(mtl_text, str) else bytes(mtl_text)
    n = C.c_size_t(0)
    ln = C.c_size_t(0)
    check(lib().er_load_obj(ob, C.c_size_t(len(ob)), mt, C.c_size_t(len(mt)), int(bool(recompute_normals)), None, C.c_size_t(0),
                            C.byref(n), None, None, C.c_size_t(0), C.byref(ln)))
    tris = np.zeros((max(n.value, 1), TRI_WORDS), dtype=np.float32)
    names = C.create_string_buffer(ln.value + 1)
    check(lib().er_load_obj(ob, C.c_size_t(len(ob)), mt, C.c_size_t(len(mt)), int(bool(recompute_normals)), ptr(tris),
                            C.c_size_t(n.value), C.byref(n), None, names, C.c_size_t(ln.value + 1), C.byref(ln)))
    return tris[:n.value], names.value.decode().split("\n")[:n.value]
