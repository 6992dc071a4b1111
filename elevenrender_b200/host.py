"""Host-side mirror of the reference's scene model and render manager for the hot path.

Names and argument meaning follow the reference (src/Scene.h, src/Camera.h, src/Material.h,
src/kernel.h:56-75 RenderParameters, src/Managers.h:41-67 RenderingManager) so tests read like the
reference's own call sites:

    scene = Scene(x_res, y_res); scene.addMaterial(...); scene.addMeshObject(tris) ...
    rm = RenderingManager(RenderParameters(w, h, sample_target, device=...))
    rm.start_rendering(scene); rm.get_render_info(); rm.get_pass("beauty")

Everything numeric goes through the C ABI (capi.py -> liber_cuda.so); nothing here computes pixels.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import (CAM_WORDS, ER_EXT_MIS, ER_EXT_POINT_LIGHTS, ER_EXT_ROULETTE, ER_PASSES, ER_RNG_LTR, ER_RNG_RTL, HIT_DTYPE, MAT_WORDS, NODE_DTYPE, TRI_WORDS,
                   ErCounters, ErRenderOpts, ErSceneDesc, ErTexture, check, f32, lib, ptr)


def camera_words(position=(0, 0, 0), rotation=(0, 0, 0), focalLength=35 * 0.001, sensorWidth=36 * 0.001,
                 sensorHeight=24 * 0.001, aperture=2.8, focusDistance=1000000.0, bokeh=False):
    """src/Camera.h:9-21 (defaults included) as the 12 ErCamera words."""
    w = np.zeros(CAM_WORDS, dtype=np.float32)
    w[0:5] = [focalLength, sensorWidth, sensorHeight, aperture, focusDistance]
    w[5:8] = rotation
    w[8:9].view(np.int32)[0] = 1 if bokeh else 0
    w[9:12] = position
    return w


_MAT_SCALARS = ["opacity", "roughness", "metallic", "clearcoatGloss", "clearcoat", "anisotropic", "eta", "transmission",
                "specular", "specularTint", "sheenTint", "subsurface", "sheen", "ax", "ay"]
_MAT_IDS = ["albedoTextureID", "emissionTextureID", "roughnessTextureID", "metallicTextureID", "normalTextureID",
            "opacityTextureID", "transmissionTextureID", "albedoShaderID"]
_MAT_DEFAULTS = dict(opacity=1, roughness=1, metallic=0, clearcoatGloss=0, clearcoat=0, anisotropic=0, eta=0,
                     transmission=0, specular=0.5, specularTint=0, sheenTint=0.5, subsurface=0, sheen=0, ax=0, ay=0)


def material_words(albedo=(0.5, 0.5, 0.5), emission=(0, 0, 0), **kw):
    """src/Material.h:21-47 (defaults included) as the 29 ErMaterial words."""
    w = np.zeros(MAT_WORDS, dtype=np.float32)
    ids = w[0:8].view(np.int32)
    for i, n in enumerate(_MAT_IDS):
        ids[i] = int(kw.pop(n, -1))
    w[8:11] = albedo
    w[11:14] = emission
    for i, n in enumerate(_MAT_SCALARS):
        w[14 + i] = kw.pop(n, _MAT_DEFAULTS[n])
    if kw:
        raise TypeError(f"unknown material fields: {sorted(kw)}")
    return w


class Texture:
    """src/Texture.h:18-28 — interleaved float texels, row-major; TCP textures are NO_FILTER."""

    def __init__(self, name, width, height, channels, data, bilinear=False):
        self.name, self.width, self.height, self.channels = name, int(width), int(height), int(channels)
        self.data = f32(data).reshape(-1)
        assert self.data.size == self.width * self.height * self.channels
        self.bilinear = bool(bilinear)

    def record(self):
        return ErTexture(self.width, self.height, self.channels, 1 if self.bilinear else 0, self.data.ctypes.data)


class HDRI:
    """src/HDRI.h — environment texture + CDF (built by er_hdri_build_cdf = HDRI::generateCDF)."""

    def __init__(self, texture=None, color=(0.5, 0.5, 0.5)):
        if texture is None:  # HDRI(Vector3 color): 1x1, 3 channels (Texture.cpp:155-162)
            texture = Texture("hdri", 1, 1, 3, np.array(color, dtype=np.float32))
        self.texture = texture
        self.cdf, self.radianceSum = capi.hdri_build_cdf(texture.data, texture.width, texture.height, texture.channels)


class Scene:
    """src/Scene.h — materials[0] is the default material (Scene.h:43)."""

    def __init__(self, x_res, y_res):
        self.x_res, self.y_res = int(x_res), int(y_res)
        self.materials = [material_words()]
        self.material_names = ["default"]
        self.textures = []
        self.tris = np.zeros((0, TRI_WORDS), dtype=np.float32)
        self.pointLights = []
        self.camera = camera_words()
        self.hdri = None
        self.meshObjectCount = 0
        self.nodes = None
        self.triIndices = None

    def addMaterial(self, words, name=""):
        self.materials.append(f32(words))
        self.material_names.append(name)
        return len(self.materials) - 1

    def addTexture(self, texture):
        self.textures.append(texture)
        return len(self.textures) - 1

    def addHDRI(self, hdri):
        self.hdri = hdri

    def addPointLight(self, position, radiance):
        self.pointLights.append((tuple(position), tuple(radiance)))  # uploaded, inert

    def addMeshObject(self, tri_words, keep_object_id=False):
        """Scene::addMeshObject (Scene.cpp:48-66): appends triangles and stamps objectID."""
        t = f32(tri_words).reshape(-1, TRI_WORDS).copy()
        if not keep_object_id:
            t[:, 37].view(np.int32)[:] = self.meshObjectCount
        self.meshObjectCount += 1
        self.tris = np.concatenate([self.tris, t], axis=0)
        self.nodes = None

    def triCount(self):
        return self.tris.shape[0]

    def buildBVH(self, nthreads=0):
        """Scene::buildBVH (Scene.cpp:122-143) through er_bvh_build."""
        self.nodes, self.triIndices = capi.bvh_build(self.tris, nthreads)
        return self.nodes, self.triIndices

    def set_bvh(self, nodes, tri_indices):
        self.nodes = np.ascontiguousarray(nodes).view(NODE_DTYPE).reshape(-1)
        self.triIndices = np.ascontiguousarray(tri_indices, dtype=np.int32)


def scene_desc(scene):
    """ErSceneDesc over the scene's host arrays (what copy_scene reads, src/SYCLCopy.cpp:3-104).
    Returns (desc, keepalive): the arrays must outlive the call that consumes the descriptor."""
    if scene.nodes is None:
        scene.buildBVH()
    if scene.hdri is None:
        scene.addHDRI(HDRI())
    d = ErSceneDesc()
    d.x_res, d.y_res = scene.x_res, scene.y_res
    tris = np.ascontiguousarray(scene.tris, dtype=np.float32)
    d.tris, d.tri_count, d.tri_stride = tris.ctypes.data, tris.shape[0], TRI_WORDS * 4
    nodes = np.ascontiguousarray(scene.nodes)
    d.nodes, d.node_stride = nodes.ctypes.data, 44
    ti = np.ascontiguousarray(scene.triIndices, dtype=np.int32)
    if ti.size == 0:
        ti = np.zeros(1, dtype=np.int32)
    d.tri_indices = ti.ctypes.data
    mats = np.ascontiguousarray(np.stack(scene.materials), dtype=np.float32)
    d.materials, d.material_count = mats.ctypes.data, mats.shape[0]
    recs = (ErTexture * max(1, len(scene.textures)))(*[t.record() for t in scene.textures])
    d.textures, d.texture_count = recs, len(scene.textures)
    d.hdri = scene.hdri.texture.record()
    cdf = f32(scene.hdri.cdf)
    d.hdri_cdf, d.hdri_radiance_sum = cdf.ctypes.data, scene.hdri.radianceSum
    cam = f32(scene.camera)
    d.camera = cam.ctypes.data
    pl = f32(np.array(scene.pointLights, dtype=np.float32).reshape(-1, 6)) if scene.pointLights else None
    d.point_lights, d.point_light_count = (pl.ctypes.data if pl is not None else None), len(scene.pointLights)
    return d, [tris, nodes, ti, mats, recs, cdf, cam, pl, scene]


def flatten_roundtrip(scene):
    """er_flatten_roundtrip: (nodes recovered from the device pair records, per-slot ErTri words, max slack)"""
    d, keep = scene_desc(scene)
    nodes = np.zeros(capi.ER_BVH_NODE_COUNT, dtype=NODE_DTYPE)
    slots = np.zeros((scene.triCount(), TRI_WORDS), dtype=np.float32)
    slack = C.c_float(0)
    check(lib().er_flatten_roundtrip(C.byref(d), ptr(nodes), ptr(slots), C.byref(slack)))
    del keep
    return nodes, slots, float(slack.value)


class RenderParameters:
    """src/kernel.h:56-75; block_size is accepted and ignored (launch shape is the library's)."""

    def __init__(self, width=1280, height=720, sampleTarget=100, denoise=False, device=0, block_size=8, max_depth=5,
                 rng_order=ER_RNG_LTR, spp_per_launch=0, kernel=0, count_stats=False, extensions=0):
        self.width, self.height, self.sampleTarget = width, height, sampleTarget
        self.denoise, self.device, self.block_size = denoise, device, block_size
        self.max_depth, self.rng_order, self.spp_per_launch = max_depth, rng_order, spp_per_launch
        self.kernel, self.count_stats = kernel, count_stats
        self.extensions = int(extensions)  # ER_EXT_* bits (off by default: they change the image)

    def opts(self):
        o = ErRenderOpts()
        o.max_depth, o.rng_order, o.spp_per_launch = self.max_depth, self.rng_order, self.spp_per_launch
        o.kernel, o.count_stats = self.kernel, 1 if self.count_stats else 0
        o.extensions = self.extensions
        return o


class RenderingManager:
    """src/Managers.h:41-67 over the C ABI: start_rendering / get_pass / get_render_info."""

    def __init__(self, pars=None, shard=(0, 1, 8), stream=None, ctx=None):
        self.pars = pars or RenderParameters()
        self.owns_ctx = ctx is None
        self.ctx = C.c_void_p() if ctx is None else ctx
        if ctx is None:
            check(lib().er_create(int(self.pars.device), C.byref(self.ctx)))
        self.shard = shard
        if shard != (0, 1, 8):
            check(lib().er_set_shard(self.ctx, *[int(v) for v in shard]))
        if stream is not None:
            check(lib().er_set_stream(self.ctx, C.c_void_p(int(stream))))
        self.scene = None
        self.P = 0

    def close(self):
        if self.ctx and self.owns_ctx:
            lib().er_destroy(self.ctx)
        self.ctx = C.c_void_p()

    def set_shard(self, index, count, rows_per_block=8):
        """interleaved row blocks (er_set_shard); restarts accumulation"""
        check(lib().er_set_shard(self.ctx, int(index), int(count), int(rows_per_block)))

    def set_stripe(self, y_begin, y_end, rows_per_block=8):
        """contiguous rows [y_begin, y_end) (er_set_shard_stripe); restarts accumulation"""
        check(lib().er_set_shard_stripe(self.ctx, int(y_begin), int(y_end), int(rows_per_block)))

    def set_block_order(self, order):
        """er_set_block_order: hand out the owned row blocks in this order (None: top to bottom)"""
        if order is None:
            check(lib().er_set_block_order(self.ctx, None, 0))
        else:
            o = np.ascontiguousarray(order, dtype=np.int32)
            check(lib().er_set_block_order(self.ctx, ptr(o), int(o.size)))

    def row_cost(self, block_first=0, block_stride=4):
        """er_row_cost: per-row work of a 1-spp counting launch over every block_stride-th row block"""
        cost = np.zeros(self.scene.y_res, dtype=np.uint32)
        o = self.pars.opts()
        check(lib().er_row_cost(self.ctx, int(block_first), int(block_stride), C.byref(o), ptr(cost)))
        return cost

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- renderSetup: copy_scene + setupKernel -------------------------------------------------
    def upload(self, scene):
        d, keep = scene_desc(scene)
        check(lib().er_upload_scene(self.ctx, C.byref(d)))
        del keep
        self.scene = scene
        self.P = scene.x_res * scene.y_res

    def set_camera(self, words):
        w = f32(words)
        check(lib().er_set_camera(self.ctx, ptr(w)))

    def reset(self):
        check(lib().er_render_reset(self.ctx))

    # -- kernel_render_enqueue --------------------------------------------------------------------
    def start_rendering(self, scene=None, blocking=False):
        if scene is not None:
            self.upload(scene)
        o = self.pars.opts()
        if blocking:
            check(lib().er_render_samples(self.ctx, int(self.pars.sampleTarget), C.byref(o)))
        else:
            check(lib().er_render_start(self.ctx, int(self.pars.sampleTarget), C.byref(o)))

    def render_samples(self, n):
        o = self.pars.opts()
        check(lib().er_render_samples(self.ctx, int(n), C.byref(o)))

    def render_enqueue(self, n):
        o = self.pars.opts()
        check(lib().er_render_enqueue(self.ctx, int(n), C.byref(o)))

    def sync(self):
        check(lib().er_sync(self.ctx))

    def wait(self):
        check(lib().er_render_wait(self.ctx))

    def abort(self):
        check(lib().er_render_abort(self.ctx))

    def poll(self):
        s, r = C.c_uint(0), C.c_int(0)
        check(lib().er_render_poll(self.ctx, C.byref(s), C.byref(r)))
        return s.value, bool(r.value)

    # -- readout ----------------------------------------------------------------------------------
    def get_render_info(self):
        s = C.c_uint(0)
        check(lib().er_get_render_info(self.ctx, C.byref(s)))
        return {"samples": s.value}

    def get_pass(self, name, out=None):
        p = ER_PASSES[name.lower()] if isinstance(name, str) else int(name)
        if out is None:
            out = np.empty((self.scene.y_res, self.scene.x_res, 4), dtype=np.float32)
        check(lib().er_get_pass(self.ctx, p, ptr(out)))
        return out

    def get_pass_owned(self, name, frame):
        """er_get_pass_owned: this shard's rows into the caller's full (h, w, 4) frame"""
        p = ER_PASSES[name.lower()] if isinstance(name, str) else int(name)
        assert frame.dtype == np.float32 and frame.size == self.P * 4
        check(lib().er_get_pass_owned(self.ctx, p, C.c_void_p(frame.ctypes.data)))
        return frame

    def get_pass_device(self, name):
        p = ER_PASSES[name.lower()] if isinstance(name, str) else int(name)
        d = C.c_void_p()
        check(lib().er_get_pass_device(self.ctx, p, C.byref(d)))
        return d.value

    def get_pixel_state(self):
        rng = np.empty(self.P, dtype=np.uint32)
        sa = np.empty(self.P, dtype=np.uint32)
        check(lib().er_get_pixel_state(self.ctx, ptr(rng), ptr(sa)))
        return rng, sa

    def get_counters(self):
        c = ErCounters()
        check(lib().er_get_counters(self.ctx, C.byref(c)))
        return c.as_dict()

    def reset_counters(self):
        check(lib().er_reset_counters(self.ctx))

    def kernel_time(self):
        ms, n = C.c_float(0), C.c_int(0)
        check(lib().er_get_kernel_time(self.ctx, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # -- nearest-hit queries ----------------------------------------------------------------------
    def trace(self, rays, mode=0):
        rays = f32(rays).reshape(-1, 6)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        check(lib().er_trace_mode(self.ctx, ptr(rays), C.c_size_t(rays.shape[0]), ptr(hits), int(mode)))
        return hits

    def device_geometry(self, which):
        """er_test_geometry: the uploaded array (0 pair records, 1 vertices, 2 normals, 3 attributes) as float32"""
        total = C.c_size_t(0)
        check(lib().er_test_geometry(self.ctx, int(which), None, C.c_size_t(0), C.byref(total)))
        out = np.zeros(total.value, dtype=np.float32)
        check(lib().er_test_geometry(self.ctx, int(which), ptr(out), C.c_size_t(out.size), None))
        return out

    def trace_noalloc(self, rays):
        """er_trace_noalloc: the same query through the library's fixed device staging (no device allocation; at most
        capi.TRACE_NOALLOC_MAX rays per call). No counterpart in the reference."""
        rays = f32(rays).reshape(-1, 6)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        check(lib().er_trace_noalloc(self.ctx, ptr(rays), C.c_size_t(rays.shape[0]), ptr(hits)))
        return hits

    def trace_device(self, d_rays, n, d_hits):
        check(lib().er_trace_device(self.ctx, C.c_void_p(int(d_rays)), C.c_size_t(int(n)), C.c_void_p(int(d_hits))))

    def camera_rays(self, xy, r5):
        xy = np.ascontiguousarray(xy, dtype=np.int32).reshape(-1, 2)
        r5 = f32(r5).reshape(-1, 5)
        out = np.zeros((xy.shape[0], 6), dtype=np.float32)
        check(lib().er_camera_rays(self.ctx, ptr(xy), ptr(r5), C.c_size_t(xy.shape[0]), ptr(out)))
        return out


class RenderingGroup:
    """RenderingManager over several GPUs of one process (er_group_*): the frame is sharded over the devices,
    the scene replicated, get_pass gathers every member's own rows over its own PCIe link."""

    def __init__(self, pars, devices):
        self.pars = pars
        self.devices = [int(d) for d in devices]
        self.g = C.c_void_p()
        ids = (C.c_int * len(self.devices))(*self.devices)
        check(lib().er_group_create(ids, len(self.devices), C.byref(self.g)))
        self.scene = None

    def close(self):
        if self.g:
            lib().er_group_destroy(self.g)
            self.g = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def member(self, i):
        ctx = C.c_void_p()
        check(lib().er_group_ctx(self.g, int(i), C.byref(ctx)))
        rm = RenderingManager(self.pars, ctx=ctx)
        rm.scene, rm.P = self.scene, self.scene.x_res * self.scene.y_res if self.scene else 0
        return rm

    def upload(self, scene):
        d, keep = scene_desc(scene)
        check(lib().er_group_upload_scene(self.g, C.byref(d)))
        del keep
        self.scene = scene

    def balance(self):
        """cost-balanced contiguous stripes (er_group_balance); returns the n+1 row boundaries"""
        yb = np.zeros(len(self.devices) + 1, dtype=np.int32)
        o = self.pars.opts()
        check(lib().er_group_balance(self.g, C.byref(o), ptr(yb)))
        return yb

    def rebalance(self):
        """feedback step (er_group_rebalance): new stripes from the kernel time every GPU spent on what was just
        rendered; restarts accumulation. Returns the n+1 row boundaries."""
        yb = np.zeros(len(self.devices) + 1, dtype=np.int32)
        check(lib().er_group_rebalance(self.g, ptr(yb)))
        return yb

    def reset(self):
        check(lib().er_group_render_reset(self.g))

    def render_samples(self, n):
        o = self.pars.opts()
        check(lib().er_group_render_samples(self.g, int(n), C.byref(o)))

    def start_rendering(self, scene=None):
        if scene is not None:
            self.upload(scene)
        o = self.pars.opts()
        check(lib().er_group_render_start(self.g, int(self.pars.sampleTarget), C.byref(o)))

    def wait(self):
        check(lib().er_group_render_wait(self.g))

    def abort(self):
        check(lib().er_group_render_abort(self.g))

    def poll(self):
        s, r = C.c_uint(0), C.c_int(0)
        check(lib().er_group_render_poll(self.g, C.byref(s), C.byref(r)))
        return s.value, bool(r.value)

    def get_render_info(self):
        s = C.c_uint(0)
        check(lib().er_group_get_render_info(self.g, C.byref(s)))
        return {"samples": s.value}

    def get_pass(self, name, out=None):
        p = ER_PASSES[name.lower()] if isinstance(name, str) else int(name)
        if out is None:
            out = np.empty((self.scene.y_res, self.scene.x_res, 4), dtype=np.float32)
        check(lib().er_group_get_pass(self.g, p, ptr(out)))
        return out
