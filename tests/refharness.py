"""TEST INFRASTRUCTURE: ctypes wrapper of oracle/_ref/liber_ref.so — the reference's own kernel-side
sources compiled for the host (oracle/Makefile, oracle/ref_harness.cpp).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_ref", "liber_ref.so")
TRI_WORDS, MAT_WORDS, CAM_WORDS, NODE_WORDS, HIT_WORDS = 39, 29, 12, 11, 21
NODE_COUNT = (2 << 18) - 1

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_scene_create.restype = C.c_void_p
        L.ref_scene_add_material.restype = C.c_int
        L.ref_scene_add_texture.restype = C.c_int
        L.ref_scene_load_obj.restype = C.c_long
        L.ref_scene_tri_count.restype = C.c_long
        L.ref_bvh_node_count.restype = C.c_long
        L.ref_hardware_threads.restype = C.c_int
        L.ref_box_intersect.restype = C.c_int
        L.ref_tri_hit.restype = C.c_int
        _lib = L
    return _lib


class _quiet:
    """the reference builder printf()s one progress line per leaf (BVH.cpp:247-249): mute fd 1"""

    def __enter__(self):
        import sys
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *a):
        lib().ref_flush_stdout()
        os.dup2(self.saved, 1)
        os.close(self.saved)
        os.close(self.null)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def sizeof_table():
    out = np.zeros(11, dtype=np.int32)
    lib().ref_sizeof_table(_p(out))
    names = ["Tri", "Node", "BVH", "Hit", "HitData", "Material", "Texture", "HDRI", "Camera", "dev_Scene", "Rng"]
    return dict(zip(names, [int(v) for v in out]))


def hardware_threads():
    return int(lib().ref_hardware_threads())


class RefScene:
    """A reference `Scene` + its deep copy, driven per pixel index."""

    def __init__(self, x_res, y_res):
        self.h = C.c_void_p(lib().ref_scene_create(int(x_res), int(y_res)))
        self.x_res, self.y_res = int(x_res), int(y_res)
        self.P = self.x_res * self.y_res
        self.finalized = False

    def close(self):
        if self.h:
            lib().ref_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_camera(self, words):
        lib().ref_scene_set_camera(self.h, _p(f32(words)))

    def add_material(self, words, name=""):
        return lib().ref_scene_add_material(self.h, _p(f32(words)), name.encode())

    def set_material(self, i, words):
        lib().ref_scene_set_material(self.h, int(i), _p(f32(words)))

    def add_texture(self, w, h, ch, data, bilinear=False, name="tex"):
        return lib().ref_scene_add_texture(self.h, int(w), int(h), int(ch), int(bool(bilinear)), _p(f32(data)),
                                           name.encode())

    def set_hdri(self, w, h, ch, data, bilinear=False):
        lib().ref_scene_set_hdri(self.h, int(w), int(h), int(ch), int(bool(bilinear)), _p(f32(data)))

    def set_hdri_color(self, r, g, b):
        lib().ref_scene_set_hdri_color(self.h, C.c_float(r), C.c_float(g), C.c_float(b))

    def add_point_light(self, pos, rad):
        lib().ref_scene_add_point_light(self.h, _p(f32(pos)), _p(f32(rad)))

    def add_tris(self, tri_words):
        t = f32(tri_words).reshape(-1, TRI_WORDS)
        lib().ref_scene_add_tris(self.h, _p(t), C.c_long(t.shape[0]))

    def load_obj(self, obj_text, mtl_text="", recompute_normals=False):
        return int(lib().ref_scene_load_obj(self.h, obj_text.encode(), mtl_text.encode(), int(recompute_normals)))

    def tri_count(self):
        return int(lib().ref_scene_tri_count(self.h))

    def get_tris(self):
        out = np.zeros((self.tri_count(), TRI_WORDS), dtype=np.float32)
        lib().ref_scene_get_tris(self.h, _p(out))
        return out

    def finalize(self, nodes=None, tri_indices=None):
        """builds the BVH with the reference builder unless a tree is supplied"""
        if nodes is not None:
            nodes = np.ascontiguousarray(nodes)
            tri_indices = np.ascontiguousarray(tri_indices, dtype=np.int32)
            assert nodes.nbytes == 44 * NODE_COUNT
            lib().ref_scene_finalize(self.h, _p(nodes), _p(tri_indices))
        else:
            with _quiet():
                lib().ref_scene_finalize(self.h, None, None)
        self.finalized = True

    def setup(self):
        lib().ref_setup(self.h)

    def get_bvh(self):
        nodes = np.zeros((NODE_COUNT, NODE_WORDS), dtype=np.float32)
        idx = np.zeros(max(1, self.tri_count()), dtype=np.int32)
        lib().ref_get_bvh(self.h, _p(nodes), _p(idx))
        return nodes, idx[:self.tri_count()]

    def get_hdri(self, want_cdf=True, want_texels=False):
        whc = np.zeros(3, dtype=np.int32)
        s = C.c_float(0)
        lib().ref_get_hdri(self.h, _p(whc), C.byref(s), None, None)
        n = int(whc[0]) * int(whc[1])
        cdf = np.zeros(n + 1, dtype=np.float32) if want_cdf else None
        tex = np.zeros(n * int(whc[2]), dtype=np.float32) if want_texels else None
        lib().ref_get_hdri(self.h, _p(whc), C.byref(s), _p(cdf), _p(tex))
        return dict(width=int(whc[0]), height=int(whc[1]), channels=int(whc[2]), radianceSum=float(s.value), cdf=cdf,
                    texels=tex)

    def render(self, passes, nthreads=0, idx_begin=0, idx_end=None):
        if nthreads <= 0:
            nthreads = hardware_threads()
        lib().ref_render(self.h, int(passes), int(nthreads), C.c_long(idx_begin),
                         C.c_long(self.P if idx_end is None else idx_end))

    def get_passes(self):
        out = np.zeros((5, self.y_res, self.x_res, 4), dtype=np.float32)
        lib().ref_get_passes(self.h, _p(out))
        return out

    def get_samples(self):
        out = np.zeros(self.P, dtype=np.uint32)
        lib().ref_get_samples(self.h, _p(out))
        return out

    def get_rng(self):
        out = np.zeros(self.P, dtype=np.uint32)
        lib().ref_get_rng(self.h, _p(out))
        return out

    def trace(self, rays, nthreads=0):
        if nthreads <= 0:
            nthreads = hardware_threads()
        rays = f32(rays).reshape(-1, 6)
        out = np.zeros((rays.shape[0], HIT_WORDS), dtype=np.float32)
        lib().ref_trace(self.h, _p(rays), C.c_long(rays.shape[0]), _p(out), int(nthreads))
        return out

    def camera_ray(self, x, y, r5):
        out = np.zeros(6, dtype=np.float32)
        lib().ref_camera_ray(self.h, int(x), int(y), _p(f32(r5)), _p(out))
        return out

    def texture_lookup(self, tex, mode, ab):
        ab = f32(ab).reshape(-1, 2)
        out = np.zeros((ab.shape[0], 3), dtype=np.float32)
        lib().ref_texture_lookup(self.h, int(tex), int(mode), _p(ab), C.c_long(ab.shape[0]), _p(out))
        return out

    def hdri_sample(self, r):
        r = f32(r).reshape(-1)
        out = np.zeros((r.size, 10), dtype=np.float32)
        pdf = np.zeros(r.size, dtype=np.float32)
        lib().ref_hdri_sample(self.h, _p(r), C.c_long(r.size), _p(out), _p(pdf))
        return out, pdf

    def generate_hitdata(self, hit_words):
        out = np.zeros(37, dtype=np.float32)
        lib().ref_generate_hitdata(self.h, _p(f32(hit_words)), _p(out))
        return out


def rng_stream(idx, n):
    st = np.zeros(n, dtype=np.uint32)
    v = np.zeros(n, dtype=np.float32)
    lib().ref_rng_stream(int(idx), int(n), _p(st), _p(v))
    return st, v


def box_intersect(ray6, b1, b2):
    return int(lib().ref_box_intersect(_p(f32(ray6)), _p(f32(b1)), _p(f32(b2))))


def tri_hit(tri_words, ray6):
    out = np.zeros(HIT_WORDS, dtype=np.float32)
    ok = lib().ref_tri_hit(_p(f32(tri_words)), _p(f32(ray6)), _p(out))
    return int(ok), out


def libm(fn, a, b=None):
    """the reference's libm (sycl::sin ... = the host's glibc): fn 0 sin, 1 cos, 2 acos, 3 atan2(a,b), 4 pow(a,b), 5 log"""
    a = f32(a).reshape(-1)
    b = None if b is None else f32(b).reshape(-1)
    out = np.zeros(a.size, dtype=np.float32)
    lib().ref_libm(int(fn), _p(a), _p(b) if b is not None else None, C.c_long(a.size), _p(out))
    return out


def spherical_mapping(dirs):
    d = f32(dirs).reshape(-1, 3)
    out = np.zeros((d.shape[0], 2), dtype=np.float32)
    lib().ref_spherical_mapping(_p(d), C.c_long(d.shape[0]), _p(out))
    return out


def disney(hitdata37, V, L, r3):
    s, e, p = np.zeros(3, np.float32), np.zeros(3, np.float32), C.c_float(0)
    lib().ref_disney(_p(f32(hitdata37)), _p(f32(V)), _p(f32(L)), _p(f32(r3)), _p(s), _p(e), C.byref(p))
    return s, e, float(p.value)


def asl_shade(i):
    rgb = np.zeros(3, np.float32)
    lib().ref_asl_shade(int(i), _p(rgb))
    return rgb
