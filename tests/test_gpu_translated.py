"""(-m gpu) parity far from the world origin: the scenes of the golden set translated by ~1e3 units. The exact hit
distance |o + d t - o| then carries rounding of the size of ulp(|o|) (6e-5 at 1e3), larger than the fixed part of
the traversal's prune margin; the margin therefore follows the magnitude of the ray origin (er_wavefront.cuh
st_step / wf_descend, er_trace.cuh er_traverse). Primary AND short secondary rays must return the triangle the
reference returns (the whole Hit record bit-identical), and the fixed-seed radiance must stay within 1e-3."""
import numpy as np
import pytest

import common
from elevenrender_b200 import host, scenes
from test_gpu_render import check_passes, render

pytestmark = pytest.mark.gpu

SHIFT = np.float32([1000.0, -700.0, 400.0])


def _translated(name):
    spec = common.make_spec(name)
    tris = spec["tris"].copy()
    tris[:, 0:9] = (tris[:, 0:9].reshape(-1, 3, 3) + SHIFT).reshape(-1, 9)
    spec["tris"] = tris
    cam = dict(spec["camera"])
    cam["position"] = tuple(np.float32(cam.get("position", (0, 0, 0))) + SHIFT)
    spec["camera"] = cam
    return spec


def _cmp(got_struct, want_words):
    got = common.hits_struct_to_words(got_struct)
    want = want_words.view(np.uint32).reshape(-1, 21)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} of {got.shape[0]} hits differ; first {bad[:5]}: {got[bad[:2]]} vs {want[bad[:2]]}"


@pytest.mark.parametrize("name", ["grid", "cornell", "spheres"])
def test_hits_translated(gpu, ref, name):
    spec = _translated(name)
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"]))
    rm.upload(common.to_ours(spec))
    rs = common.to_ref(spec)
    v = spec["tris"][:, 0:9].reshape(-1, 3)
    lo, hi = v.min(axis=0) - 0.5, v.max(axis=0) + 0.5
    rays = scenes.random_rays(20000, origin_box=(lo, hi), target_box=(lo, hi), seed=33)
    first = rs.trace(rays)
    _cmp(rm.trace(rays, mode=0), first)
    # secondary rays: from the hit points of the first set, in random directions (short, grazing, self-intersecting)
    fw = np.asarray(first).view(np.float32).reshape(-1, 21)
    valid = np.asarray(first).view(np.uint32).reshape(-1, 21)[:, 15] != 0
    rng = np.random.default_rng(34)
    d = rng.normal(size=(int(valid.sum()), 3)).astype(np.float32)
    sec = np.concatenate([fw[valid, 0:3], d], axis=1).astype(np.float32)
    assert sec.shape[0] > 1000
    _cmp(rm.trace(sec, mode=0), rs.trace(sec))
    _cmp(rm.trace(sec, mode=2), rs.trace(sec))
    rm.close()
    rs.close()


@pytest.mark.parametrize("name", ["grid", "cornell"])
@pytest.mark.parametrize("kernel", [0, 1])
def test_radiance_translated(gpu, ref, name, kernel):
    spec = _translated(name)
    rs = common.to_ref(spec)
    rs.render(4)
    want = rs.get_passes()[[0, 2, 3, 4]]
    rs.close()
    passes, _, _, _, _ = render(spec, 4, kernel=kernel)
    check_passes(passes, want, f"{name} translated kernel={kernel}")
