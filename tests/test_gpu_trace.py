"""T3 (-m gpu): nearest-hit parity. Same BVH, same host-generated ray set on both sides; the whole
84-byte Hit record (valid, triIdx, tu, tv, position, normals ...) must be bit-identical."""
import os

import numpy as np
import pytest

import common
from elevenrender_b200 import host, scenes

pytestmark = pytest.mark.gpu

SCENES = ["grid", "grid_rot", "cornell", "textured", "spheres", "miss"]


def _cmp(got_struct, want_words):
    got = common.hits_struct_to_words(got_struct)
    want = want_words.view(np.uint32).reshape(-1, 21)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} of {got.shape[0]} hits differ; first {bad[:5]}: {got[bad[:2]]} vs {want[bad[:2]]}"


@pytest.mark.parametrize("name", SCENES)
def test_hits_match_golden(gpu, name):
    g = np.load(os.path.join(common.GOLDEN_DIR, f"{name}.npz"))
    spec = common.make_spec(name)
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"]))
    rm.upload(common.to_ours(spec))
    rays = common.primary_rays(spec)
    _cmp(rm.trace(rays, mode=0), g["hits"])   # ordered + pruned traversal (the product path)
    _cmp(rm.trace(rays, mode=2), g["hits"])   # reference visit order, no pruning
    rm.close()


@pytest.mark.parametrize("name", ["grid", "spheres", "cornell"])
def test_hits_match_oracle_random_rays(gpu, ref, name):
    """incoherent rays from random origins (inside and outside the geometry), axis-parallel rays,
    rays starting on box slabs, zero-length directions"""
    spec = common.make_spec(name)
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"]))
    sc = common.to_ours(spec)
    rm.upload(sc)
    rs = common.to_ref(spec)
    lo = spec["tris"][:, 0:9].reshape(-1, 3).min(axis=0) - 0.5
    hi = spec["tris"][:, 0:9].reshape(-1, 3).max(axis=0) + 0.5
    rays = scenes.random_rays(20000, origin_box=(lo, hi), target_box=(lo, hi), seed=21)
    rays[:300, 3:] = np.float32([0, 0, 1])
    rays[300:600, 3:] = np.float32([1, 0, 0])
    rays[600:900, 3:] = np.float32([0, -1, 0])
    rays[900:910, 3:] = 0.0
    rays[910:1200, 0:3] = sc.nodes["b1"][1:291]  # origins exactly on child-box corners
    _cmp(rm.trace(rays, mode=0), rs.trace(rays))
    rm.close()
    rs.close()


def test_counters_and_pruning(gpu):
    """the pruned traversal visits a subset of the reference visit set and returns the same hits"""
    spec = common.make_spec("grid")
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"]))
    rm.upload(common.to_ours(spec))
    rays = common.primary_rays(spec)
    rm.reset_counters()
    a = rm.trace(rays, mode=1)
    ca = rm.get_counters()
    rm.reset_counters()
    b = rm.trace(rays, mode=3)
    cb = rm.get_counters()
    assert np.array_equal(common.hits_struct_to_words(a), common.hits_struct_to_words(b))
    assert ca["rays"] == cb["rays"] == rays.shape[0]
    assert 0 < ca["box_tests"] <= cb["box_tests"] and ca["tri_tests"] <= cb["tri_tests"]
    rm.close()


def test_primary_hits_full_size(gpu, ref):
    """BASELINE config 5: 250k-triangle mesh, 2048x2048 host-generated primary rays. The oracle traces a
    strided 1/16 sample (same rays); the CUDA path traces all of them and must agree bit for bit on the
    sample, and the ordered/pruned traversal must agree with the reference-order one on ALL rays."""
    spec = dict(common.make_spec("grid"))
    spec["tris"] = scenes.displaced_grid(354, 353)
    spec["x_res"], spec["y_res"] = 2048, 2048
    sc = common.to_ours(spec)
    sc.buildBVH()
    rm = host.RenderingManager(host.RenderParameters(2048, 2048))
    rm.upload(sc)
    rays = common.primary_rays(spec)
    got = rm.trace(rays, mode=0)
    got_ref_order = rm.trace(rays, mode=2)
    assert np.array_equal(common.hits_struct_to_words(got), common.hits_struct_to_words(got_ref_order))
    assert (got["valid"] != 0).mean() > 0.5
    rs = common.to_ref(spec, nodes=sc.nodes, tri_indices=sc.triIndices)  # builder identity is tested separately
    sel = np.arange(0, rays.shape[0], 16)
    _cmp(got[sel], rs.trace(rays[sel]))
    rm.close()
    rs.close()
