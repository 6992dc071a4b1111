"""(-m gpu) er_upload_scene with page-locked triangles: the records cross PCIe as they are and the device builds
the per-slot layout (k_flatten_tris) instead of the host (erh::flatten_geometry). Both paths must leave the same
arrays on the device bit for bit — pair records with their subtree slack, vertices, normals, attributes — and
therefore the same hits, frames and traversal counters. Also: bad indices / material ids are refused on both."""
import numpy as np
import pytest

import common
from elevenrender_b200 import capi, host, scenes

pytestmark = pytest.mark.gpu


def _scene(n=300):
    spec = common.make_spec("grid")
    spec["tris"] = scenes.displaced_grid(n, n)  # 2 n^2 triangles: above the 65536 threshold of the device path
    # a few awkward records: a degenerate triangle, a non-unit normal, a huge coordinate
    spec["tris"][7, 0:9] = spec["tris"][7, 0:3].tolist() * 3
    spec["tris"][11, 18:21] *= 3.0
    spec["tris"][13, 0] = 1e6
    spec["x_res"], spec["y_res"] = 96, 64
    return spec


def _everything(rm, rays):
    geo = [rm.device_geometry(k) for k in range(4)]
    hits = common.hits_struct_to_words(rm.trace(rays))
    rm.pars.count_stats = True
    rm.reset()
    rm.reset_counters()
    rm.render_samples(3)
    frame = rm.get_pass("beauty")
    counters = rm.get_counters()
    rm.pars.count_stats = False
    return geo, hits, frame, counters


def test_device_flatten_equals_host_flatten(gpu):
    spec = _scene()
    sc = common.to_ours(spec)
    sc.buildBVH()
    assert sc.tris.shape[0] >= 65536 and sc.tris.dtype == np.float32 and sc.tris.flags["C_CONTIGUOUS"]
    rays = common.primary_rays(spec)
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], 3, rng_order=host.ER_RNG_LTR))
    rm.upload(sc)  # pageable: the host flattens
    want = _everything(rm, rays)
    capi.host_register(sc.tris)
    try:
        rm.upload(sc)  # page-locked: the device flattens
        got = _everything(rm, rays)
        rm.upload(sc)  # and again (buffers reused)
        again = _everything(rm, rays)
    finally:
        capi.host_unregister(sc.tris)
    rm.upload(sc)  # back to the host path
    back = _everything(rm, rays)
    rm.close()
    for other, what in ((got, "device"), (again, "device, second upload"), (back, "host again")):
        for k, name in enumerate(("pair records", "vertices", "normals", "attributes")):
            assert np.array_equal(want[0][k].view(np.uint32), other[0][k].view(np.uint32)), f"{what}: {name} differ"
        assert np.array_equal(want[1], other[1]), f"{what}: hits differ"
        assert np.array_equal(want[2].view(np.uint32), other[2].view(np.uint32)), f"{what}: frame differs"
        assert want[3] == other[3], f"{what}: traversal counters differ"


def test_device_flatten_refuses_bad_records(gpu):
    spec = _scene(200)
    sc = common.to_ours(spec)
    sc.buildBVH()
    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], 1))
    for pinned in (False, True):
        if pinned:
            capi.host_register(sc.tris)
        try:
            keep = sc.tris[5, 38]
            sc.tris[5, 38] = np.int32(99).view(np.float32)  # materialID beyond the material table
            with pytest.raises(capi.ErError) as e:
                rm.upload(sc)
            assert "materialID" in str(e.value)
            sc.tris[5, 38] = keep
            old = int(sc.triIndices[9])
            sc.triIndices[9] = sc.tris.shape[0] + 4
            with pytest.raises(capi.ErError) as e:
                rm.upload(sc)
            assert "tri_indices" in str(e.value)
            sc.triIndices[9] = old
            rm.upload(sc)  # the context still works
            rm.render_samples(1)
        finally:
            if pinned:
                capi.host_unregister(sc.tris)
    rm.close()
