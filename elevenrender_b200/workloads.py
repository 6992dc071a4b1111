"""The BASELINE.json configurations as scene specs.  A spec is a plain dict:
tris (n,39) float32, materials [kwargs of host.material_words], textures [(w,h,ch,data,bilinear)],
hdri (w,h,ch,data,bilinear) | (r,g,b), camera kwargs, x_res, y_res, spp, point_lights."""
import numpy as np

from . import host, scenes


def _c4_materials():
    mats = []
    for i in range(16):
        kw = dict(albedo=(0.25 + 0.04 * i, 0.7 - 0.03 * i, 0.45), specular=0.2 + 0.04 * i, roughness=0.2 + 0.04 * i,
                  metallic=(i % 4) / 4.0, clearcoat=(i % 3) * 0.4, clearcoatGloss=(i % 5) / 5.0,
                  anisotropic=(i % 4) * 0.2, sheen=(i % 2) * 0.5, subsurface=(i % 3) * 0.25)
        kw["albedoTextureID"] = 0
        kw["roughnessTextureID"] = 1
        if i % 2:
            kw["metallicTextureID"] = 2
        if i % 4 != 3:
            kw["normalTextureID"] = 3
        mats.append(kw)
    return mats


def workload(name):
    s = dict(name=name, materials=[], textures=[], hdri=(0.5, 0.5, 0.5), point_lights=[])
    if name == "c1":
        tris, mats = scenes.cornell_box()
        s.update(tris=tris, materials=mats, camera=scenes.cornell_camera(), x_res=256, y_res=256, spp=16,
                 point_lights=[((0.0, 1.8, 0.0), (10.0, 10.0, 10.0))],
                 desc="procedural Cornell box (36 tris), 256x256, 16 spp, Disney diffuse, 1 (inert) point light")
    elif name == "c2":
        s.update(tris=scenes.displaced_grid(354, 353), camera=scenes.grid_camera(), x_res=1920, y_res=1080, spp=64,
                 hdri=(2048, 1024, 3, scenes.sky_hdri(2048, 1024), False),
                 desc="procedural 249,924-tri displaced grid, 1920x1080, 64 spp, synthetic 2048x1024 HDRI env + NEE")
    elif name == "c3":
        s.update(tris=scenes.sphere_field(), camera=scenes.sphere_camera(), x_res=1920, y_res=1080, spp=256,
                 hdri=(2048, 1024, 3, scenes.sky_hdri(2048, 1024), False),
                 desc="procedural 10,000,000-tri displaced-sphere field, 1920x1080, 256 spp")
    elif name == "c4":
        m = scenes.material_maps(1024)
        s.update(tris=scenes.displaced_grid(354, 353, unit_uv=False, material_cells=12), camera=scenes.grid_camera(),
                 x_res=3840, y_res=2160, spp=1024, materials=_c4_materials(),
                 textures=[(1024, 1024, 3, m["albedo"], False), (1024, 1024, 1, m["roughness"], False),
                           (1024, 1024, 1, m["metallic"], False), (1024, 1024, 3, m["normal"], False)],
                 hdri=(2048, 1024, 3, scenes.sky_hdri(2048, 1024), False),
                 desc="249,924-tri grid, 16 textured Disney materials (1024^2 maps), 3840x2160, 1024 spp")
    elif name == "c5":
        s.update(tris=scenes.displaced_grid(354, 353), camera=scenes.grid_camera(), x_res=2048, y_res=2048, spp=1,
                 hdri=(2048, 1024, 3, scenes.sky_hdri(2048, 1024), False),
                 desc="primary-ray hit-ID check + depth-16 stress, 2048x2048, 249,924-tri grid")
    else:
        raise KeyError(name)
    return s


def scene_from_spec(spec):
    """host.Scene (the CUDA side's view) of a spec"""
    sc = host.Scene(spec["x_res"], spec["y_res"])
    for kw in spec["materials"]:
        sc.addMaterial(host.material_words(**kw))
    for i, (w, h, ch, data, bil) in enumerate(spec["textures"]):
        sc.addTexture(host.Texture(f"tex{i}", w, h, ch, data, bilinear=bil))
    hd = spec["hdri"]
    if len(hd) == 3:
        sc.addHDRI(host.HDRI(color=hd))
    else:
        sc.addHDRI(host.HDRI(host.Texture("hdri", hd[0], hd[1], hd[2], hd[3], bilinear=hd[4])))
    for (p, r) in spec["point_lights"]:
        sc.addPointLight(p, r)
    sc.camera = host.camera_words(**spec["camera"])
    if spec["tris"].shape[0]:
        sc.addMeshObject(spec["tris"], keep_object_id=True)
    return sc
