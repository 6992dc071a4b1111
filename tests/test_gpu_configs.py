"""Parity on the BASELINE configurations AT THEIR OWN SIZE (-m gpu): every config's full scene at its own
resolution, a reduced sample count, RTL draw order, and >= 100k pixels in separated row bands compared with
the live oracle (the reference's own renderingKernel per pixel index over the same index ranges).

  C1  Cornell box through the reference's ObjLoader stream path (OBJ + MTL text), 256x256, 16 spp, whole frame
  C2  249,924-triangle grid, 2048x1024 HDRI (21-probe CDF search), 1920x1080
  C3  10,000,000-triangle sphere field (38-triangle leaves: the warp-cooperative leaf tests), 1920x1080
  C4  16 textured Disney materials with 1024^2 maps, 3840x2160
  C5  primary rays only on C3 geometry at 2048x2048, hit records bit-exact on a 1/64 ray sample

The measured outlier fraction (pixels whose radiance differs by more than 1e-3 relative) of every pass of
every config is appended to gpurun_out/parity_outliers.jsonl; DESIGN.md §5 holds the table of one run.
The bound is SURVEY T4's: at most 1e-4 of the compared pixels.
"""
import json
import os
import time

import numpy as np
import pytest

import common
from elevenrender_b200 import host, scenes, workloads

pytestmark = pytest.mark.gpu

REL_TOL = 1e-3
ABS_FLOOR = 1e-5
MAX_OUTLIERS = 1e-4


def record(config, what, fracs, pixels, extra=None):
    line = {"config": config, "what": what, "pixels": int(pixels), "rel_tol": REL_TOL,
            "outlier_fraction": {k: float(v) for k, v in zip(("beauty", "normal", "tangent", "bitangent"), fracs)}}
    line.update(extra or {})
    print("PARITY", json.dumps(line))
    d = os.path.join(common.ROOT, "gpurun_out")
    try:
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "parity_outliers.jsonl"), "a") as f:
            f.write(json.dumps(line) + "\n")
    except OSError:
        pass


def bands_for(y_res, x_res, want_pixels=100_000, n_bands=4):
    """n_bands separated row bands (top, upper middle, lower middle, bottom) holding >= want_pixels pixels"""
    rows = -(-want_pixels // (x_res * n_bands))
    rows = max(rows, 2)
    starts = [int(round((y_res - rows) * f)) for f in np.linspace(0.07, 0.93, n_bands)]
    return [(y0, min(y_res, y0 + rows)) for y0 in starts]


def compare_bands(config, spec, spp, rs, bands, max_depth=5):
    """render the full frame on the GPU, the bands on the oracle; per-pass outlier fractions over the bands"""
    x_res = spec["x_res"]
    pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_RTL, spp_per_launch=spp,
                                 max_depth=max_depth)
    rm = host.RenderingManager(pars)
    sc = common.to_ours(spec) if "scene" not in spec else spec["scene"]
    if "nodes" in spec:
        sc.set_bvh(spec["nodes"], spec["tri_indices"])
    rm.upload(sc)
    rm.render_samples(spp)
    got = np.stack([rm.get_pass(p) for p in ("beauty", "normal", "tangent", "bitangent")])
    _, sa = rm.get_pixel_state()
    rm.close()
    t0 = time.time()
    for (y0, y1) in bands:
        rs.render(spp, 0, y0 * x_res, y1 * x_res)
    t_ref = time.time() - t0
    want = rs.get_passes()[[0, 2, 3, 4]]
    want_sa = rs.get_samples().reshape(spec["y_res"], spec["x_res"])
    sel = np.zeros(spec["y_res"], dtype=bool)
    for (y0, y1) in bands:
        sel[y0:y1] = True
    g, w = got[:, sel], want[:, sel]
    err = common.rel_err(g[..., :3], w[..., :3], floor=ABS_FLOOR).max(axis=-1)
    fracs = (err > REL_TOL).mean(axis=(1, 2))
    pixels = int(sel.sum()) * x_res
    hit_frac = float((w[1, ..., :3] != 0).any(axis=-1).mean())
    bit_equal = (g[..., :3].view(np.uint32) == w[..., :3].view(np.uint32)).all(axis=-1).mean(axis=(1, 2))
    record(config, f"{spp} spp RTL, bands {bands}", fracs, pixels,
           {"oracle_s": round(t_ref, 2), "first_hit_pixel_fraction": hit_frac, "beauty_mean": float(w[0, ..., :3].mean()),
            "max_rel_err": [float(e) for e in err.max(axis=(1, 2))],
            "bit_identical_pixel_fraction": [float(b) for b in bit_equal]})
    assert np.array_equal(g[..., 3], w[..., 3]), f"{config}: alpha differs"
    assert np.array_equal(sa.reshape(spec["y_res"], spec["x_res"])[sel], want_sa[sel]), f"{config}: sample counters differ"
    assert fracs.max() <= MAX_OUTLIERS, f"{config}: outlier fraction per pass {fracs} > {MAX_OUTLIERS} (tol {REL_TOL})"
    # with the host libm's bits on the device (er_libm.cuh) the passes are the reference's bit for bit; 1e-3 of the
    # pixels are allowed to differ (the restated sinf/cosf/powf differ from glibc's FMA build in ~1e-8 of calls)
    assert bit_equal.min() >= 0.999, f"{config}: bit-identical pixel fraction per pass {bit_equal}"
    assert pixels >= 60_000
    return fracs


def test_c1_cornell_obj_full_frame(gpu, ref):
    """BASELINE config 1 as BASELINE.md §3 specifies it: OBJ text through the reference's ObjLoader stream path
    (ObjLoader.cpp:154-164: triangulate, z-flip, MikkTSpace tangents), default Disney diffuse + emissive quad,
    one (inert) point light, 256x256, 16 spp — the WHOLE frame against the oracle."""
    spec = workloads.workload("c1")
    obj, mtl = scenes.cornell_obj_text()
    rs = ref.RefScene(spec["x_res"], spec["y_res"])
    for kw, name in zip(spec["materials"], ("red", "green", "light")):
        rs.add_material(host.material_words(**kw), name)
    rs.set_hdri_color(*spec["hdri"])
    for (p, r) in spec["point_lights"]:
        rs.add_point_light(p, r)
    rs.set_camera(host.camera_words(**spec["camera"]))
    with ref._quiet():
        n = rs.load_obj(obj, mtl)
    assert n == spec["tris"].shape[0] == 36
    loaded = rs.get_tris()  # what the reference's loader made of the text (tangents from MikkTSpace)
    assert np.array_equal(loaded[:, :9], spec["tris"][:, :9]), "the loader's z flip restores cornell_box()"
    assert np.array_equal(loaded[:, 38].view(np.int32), spec["tris"][:, 38].view(np.int32)), "pair_materials by name"
    rs.finalize()
    rs.setup()
    spec = dict(spec)
    spec["tris"] = loaded
    fr = compare_bands("c1", spec, 16, rs, [(0, spec["y_res"])])
    rs.close()
    assert fr.shape == (4,)


def test_c2_full_size_bands(gpu, ref):
    spec = workloads.workload("c2")
    rs = common.to_ref(spec)  # the reference's own BVH::build (~6 s)
    compare_bands("c2", spec, 8, rs, bands_for(spec["y_res"], spec["x_res"], 150_000))
    rs.close()


def test_c3_full_size_bands_and_primary_hits(gpu, ref):
    """C3 + C5 on C3 geometry. The oracle renders over the tree of er_bvh_build, which is bit-identical to the
    reference builder's at this size (tests/test_oracle_cpu.py::test_bvh_builder_identical_c3_checksum; the
    reference's own build takes 87 s and several GB here)."""
    spec = workloads.workload("c3")
    sc = common.to_ours(spec)
    sc.buildBVH()
    rs = common.to_ref(spec, sc.nodes, sc.triIndices)
    spec = dict(spec)
    spec["scene"] = sc
    compare_bands("c3", spec, 4, rs, bands_for(spec["y_res"], spec["x_res"], 120_000))
    # C5 on C3 geometry: 2048^2 host-generated primary rays; ordered+pruned traversal == reference-order
    # traversal on ALL rays; the oracle checks a strided 1/64 sample bit for bit
    s5 = dict(spec)
    s5["x_res"], s5["y_res"] = 2048, 2048
    rays = common.primary_rays(s5)
    rm = host.RenderingManager(host.RenderParameters(64, 64))
    sc.x_res, sc.y_res = 64, 64
    rm.upload(sc)
    got = rm.trace(rays, mode=0)
    got_ref_order = rm.trace(rays, mode=2)
    assert np.array_equal(common.hits_struct_to_words(got), common.hits_struct_to_words(got_ref_order))
    sel = np.arange(17, rays.shape[0], 64)
    want = rs.trace(rays[sel]).view(np.uint32).reshape(-1, 21)
    g = common.hits_struct_to_words(got[sel])
    bad = np.nonzero((g != want).any(axis=1))[0]
    print("PARITY", json.dumps({"config": "c5/c3-geometry", "rays": int(rays.shape[0]), "oracle_rays": int(sel.size),
                                "valid_fraction": float((got["valid"] != 0).mean()), "mismatches": int(bad.size)}))
    assert bad.size == 0, f"{bad.size} of {sel.size} primary hits differ on C3 geometry"
    assert (got["valid"] != 0).mean() > 0.2
    rm.close()
    rs.close()


def test_c4_full_size_bands(gpu, ref):
    spec = workloads.workload("c4")
    rs = common.to_ref(spec)
    compare_bands("c4", spec, 4, rs, bands_for(spec["y_res"], spec["x_res"], 120_000))
    rs.close()


def test_c5_primary_hits_c2_reference_tree(gpu, ref):
    """C5 on C2 geometry with the tree of the REFERENCE's own builder on the oracle side and the tree of
    er_bvh_build on ours: hit records bit-exact on a 1/16 ray sample of the 2048^2 rays."""
    spec = workloads.workload("c5")
    rm = host.RenderingManager(host.RenderParameters(64, 64))
    sc = common.to_ours(spec)
    sc.x_res, sc.y_res = 64, 64
    rm.upload(sc)
    rays = common.primary_rays(spec)
    got = rm.trace(rays, mode=0)
    rs = common.to_ref(spec)  # BVH::build itself
    sel = np.arange(5, rays.shape[0], 16)
    want = rs.trace(rays[sel]).view(np.uint32).reshape(-1, 21)
    g = common.hits_struct_to_words(got[sel])
    assert np.array_equal(g, want)
    rm.close()
    rs.close()
