"""Shared test scenes: one spec feeds both the CUDA path (elevenrender_b200.host.Scene) and the
reference oracle (tests/refharness.RefScene), so both sides see the same bits."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from elevenrender_b200 import host, scenes  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def _textured_materials():
    """16 materials over 4 maps (albedo 0, roughness 1, metallic 2, normal 3) with the scalar Disney
    parameters swept; a few with opacity < 1, emission, and one custom-shader slot."""
    mats = []
    for i in range(16):
        kw = dict(albedo=(0.2 + 0.05 * i, 0.7 - 0.03 * i, 0.4), specular=0.1 + 0.05 * i, roughness=0.15 + 0.05 * i,
                  metallic=(i % 4) / 4.0, clearcoat=(i % 3) * 0.5, clearcoatGloss=(i % 5) / 5.0,
                  anisotropic=(i % 4) * 0.2, sheen=(i % 2) * 0.6, sheenTint=0.3, subsurface=(i % 3) * 0.3,
                  specularTint=(i % 5) * 0.2)
        if i % 2 == 0:
            kw["albedoTextureID"] = 0
        if i % 3 == 0:
            kw["roughnessTextureID"] = 1
        if i % 4 == 1:
            kw["metallicTextureID"] = 2
        if i % 4 == 2:
            kw["normalTextureID"] = 3
        if i == 5:
            kw["opacity"] = 0.6
        if i == 7:
            kw["emission"] = (0.5, 0.4, 0.3)
        if i == 9:
            kw["albedoShaderID"] = 1
        if i == 11:
            kw["transmission"] = 1.0
        mats.append(kw)
    return mats


def make_spec(name):
    """dict(tris, materials[kw], textures[(w,h,ch,data,bilinear)], hdri=(w,h,ch,data,bilinear)|color, camera kw,
    x_res, y_res, point_lights)"""
    s = dict(materials=[], textures=[], hdri=(0.5, 0.5, 0.5), point_lights=[], x_res=48, y_res=32)
    if name == "grid":
        s["tris"] = scenes.displaced_grid(40, 30)
        s["camera"] = scenes.grid_camera()
        s["hdri"] = (64, 32, 3, scenes.sky_hdri(64, 32), False)
    elif name == "grid_rot":  # rotated camera + thin lens
        s["tris"] = scenes.displaced_grid(24, 20)
        s["camera"] = dict(position=(0.4, -0.2, -1.1), rotation=(3.0, -7.0, 11.0), bokeh=True, aperture=1.4,
                           focusDistance=4.0)
        s["hdri"] = (32, 16, 3, scenes.sky_hdri(32, 16), True)
    elif name == "cornell":
        tris, mats = scenes.cornell_box()
        s["tris"], s["materials"] = tris, mats
        s["camera"] = scenes.cornell_camera()
        s["point_lights"] = [((0.0, 1.8, 0.0), (10.0, 10.0, 10.0))]
        s["x_res"], s["y_res"] = 40, 40
    elif name == "textured":
        s["tris"] = scenes.displaced_grid(32, 32, unit_uv=False, material_cells=4)
        s["materials"] = _textured_materials()
        m = scenes.material_maps(64)
        s["textures"] = [(64, 64, 3, m["albedo"], False), (64, 64, 1, m["roughness"], False),
                         (64, 64, 1, m["metallic"], True), (64, 64, 3, m["normal"], False)]
        s["camera"] = scenes.grid_camera()
        s["hdri"] = (128, 64, 3, scenes.sky_hdri(128, 64), False)
    elif name == "spheres":
        s["tris"] = scenes.sphere_field(side=3, lon=16, lat=8, spacing=0.5, radius=0.2, z0=2.0)
        s["camera"] = scenes.sphere_camera()
        s["hdri"] = (64, 32, 3, scenes.sky_hdri(64, 32), False)
    elif name == "dense":  # 1.2 M triangles: leaves hold 4-5 triangles, leaf pairs ~9 (the 10M-triangle regime in small)
        s["tris"] = scenes.sphere_field(side=5, lon=160, lat=150, spacing=0.5, radius=0.22, z0=2.2)
        s["camera"] = scenes.sphere_camera()
        s["hdri"] = (64, 32, 3, scenes.sky_hdri(64, 32), False)
    elif name == "miss":  # one triangle behind the camera: every path is an environment miss (SURVEY F3
        # known answer). The reference cannot build an empty scene (BVH.cpp:248 divides by totalTris).
        v = np.float32([[[-1, -1, -50], [1, -1, -50], [0, 1, -50]]])
        s["tris"] = scenes.pack_tris(v, np.broadcast_to(np.float32([0, 0, 1]), (1, 3, 3)))
        s["camera"] = dict(position=(0.3, 0.2, -1.0))
        s["x_res"], s["y_res"] = 16, 8
    else:
        raise KeyError(name)
    return s


def to_ours(spec):
    from elevenrender_b200 import workloads
    return workloads.scene_from_spec(spec)


def to_ref(spec, nodes=None, tri_indices=None):
    import refharness as R
    rs = R.RefScene(spec["x_res"], spec["y_res"])
    for kw in spec["materials"]:
        rs.add_material(host.material_words(**kw))
    for i, (w, h, ch, data, bil) in enumerate(spec["textures"]):
        rs.add_texture(w, h, ch, data, bilinear=bil, name=f"tex{i}")  # Scene::addTexture drops duplicate names
    hd = spec["hdri"]
    if len(hd) == 3:
        rs.set_hdri_color(*hd)
    else:
        rs.set_hdri(hd[0], hd[1], hd[2], hd[3], bilinear=hd[4])
    for (p, r) in spec["point_lights"]:
        rs.add_point_light(p, r)
    rs.set_camera(host.camera_words(**spec["camera"]))
    if spec["tris"].shape[0]:
        rs.add_tris(spec["tris"])
    rs.finalize(nodes, tri_indices)
    rs.setup()
    return rs


def primary_rays(spec, seed=7, n=None):
    """camera-ish rays generated ON THE HOST (shared by both sides): from the camera position through
    a jittered grid over the sensor, no rotation."""
    cam = host.camera_words(**spec["camera"])
    xr, yr = spec["x_res"], spec["y_res"]
    rng = np.random.default_rng(seed)
    ys, xs = np.meshgrid(np.arange(yr), np.arange(xr), indexing="ij")
    jx, jy = rng.random((yr, xr)), rng.random((yr, xr))
    sx = (-cam[1] / 2) + (xs + jx) / xr * cam[1]
    sy = (-cam[2] / 2) + (ys + jy) / yr * cam[2]
    d = np.stack([sx, sy, np.full_like(sx, cam[0])], axis=-1).reshape(-1, 3)
    o = np.broadcast_to(cam[9:12], d.shape)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    return rays if n is None else rays[:n]


def rel_err(a, b, floor=1e-5):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)


def hits_struct_to_words(h):
    """ErHit structured array -> (n,21) uint32 words as the oracle dumps them"""
    return np.ascontiguousarray(h).view(np.uint32).reshape(-1, 21)
