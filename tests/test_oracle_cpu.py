"""CPU tests (-m "not gpu"): the oracle against the committed golden vectors, and the host-side
prerequisites of the CUDA path (BVH builder, HDRI CDF, flatten round trip) against the oracle."""
import hashlib
import os

import numpy as np
import pytest

import common
from elevenrender_b200 import capi, host

SCENES = ["grid", "grid_rot", "cornell", "textured", "spheres", "miss"]


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def golden(name):
    return np.load(os.path.join(common.GOLDEN_DIR, f"{name}.npz"))


def test_sizeof_table(ref):
    # SURVEY §4 T0
    t = ref.sizeof_table()
    assert t == {"Tri": 192, "Node": 44, "BVH": 23068712, "Hit": 84, "HitData": 148, "Material": 376, "Texture": 72,
                 "HDRI": 88, "Camera": 48, "dev_Scene": 120, "Rng": 4}
    assert np.array_equal(golden("kat")["sizeof"], np.array(list(t.values())))


def test_rng_known_answers(ref):
    g = golden("kat")
    for idx in (0, 1, 12345, 2073599):
        st, v = ref.rng_stream(idx, 16)
        assert np.array_equal(st, g[f"rng_state_{idx}"])
        assert np.array_equal(v.view(np.uint32), g[f"rng_value_{idx}"].view(np.uint32))
        # restated in numpy: jenkins(idx+1) then xorshift32 13/17/5, float(state)/2^32 (kernel.cpp:25-47)
        h = 0
        seed = idx + 1
        for i in range(4):
            h = (h + ((seed >> (8 * i)) & 0xFF)) & 0xFFFFFFFF
            h = (h + (h << 10)) & 0xFFFFFFFF
            h ^= h >> 6
        h = (h + (h << 3)) & 0xFFFFFFFF
        h ^= h >> 11
        h = (h + (h << 15)) & 0xFFFFFFFF
        s = h
        for i in range(16):
            s ^= (s << 13) & 0xFFFFFFFF
            s ^= s >> 17
            s ^= (s << 5) & 0xFFFFFFFF
            assert s == int(st[i])
            assert np.float32(s) / np.float32(4294967296.0) == v[i]


def test_f3_known_answer():
    # constant 0.5 environment, every path misses -> 0.5 * N / (N + 1); samples = N + 1
    g = golden("miss")
    assert np.all(g["passes_1"][0, ..., :3] == np.float32(0.25))
    assert np.allclose(g["passes_4"][0, ..., :3], 0.4, rtol=0, atol=1e-7)
    assert np.all(g["passes_4"][0, ..., 3] == 1.0)
    assert np.all(g["samples_4"] == 5)


@pytest.mark.parametrize("name", SCENES)
def test_oracle_matches_golden(ref, name):
    """the oracle built on this machine reproduces the committed vectors bit for bit"""
    g = golden(name)
    spec = common.make_spec(name)
    rs = common.to_ref(spec)
    nodes, idx = rs.get_bvh()
    nw = nodes.view(np.uint32).reshape(-1, 11).copy()
    nw[:, 10] &= 0xFF
    assert np.array_equal(sha(nw), g["bvh_nodes_sha"])
    assert np.array_equal(sha(idx), g["bvh_idx_sha"])
    assert np.array_equal(rs.trace(common.primary_rays(spec)).view(np.uint32), g["hits"])
    rs.render(1)
    assert np.array_equal(rs.get_passes()[[0, 2, 3, 4]].view(np.uint32), g["passes_1"].view(np.uint32))
    rs.render(3)
    assert np.array_equal(rs.get_passes()[[0, 2, 3, 4]].view(np.uint32), g["passes_4"].view(np.uint32))
    assert np.array_equal(rs.get_rng(), g["rng_4"])
    assert np.array_equal(rs.get_samples(), g["samples_4"])
    rs.close()


@pytest.mark.parametrize("name", SCENES)
def test_bvh_builder_identical(name):
    """er_bvh_build reproduces BVH::build's nodes[] and triIndices[] (checked through the golden checksums,
    so this also runs where the oracle is absent)"""
    g = golden(name)
    spec = common.make_spec(name)
    for threads in (1, 4):
        nodes, idx = capi.bvh_build(spec["tris"], threads)
        nw = nodes.view(np.uint32).reshape(-1, 11).copy()
        nw[:, 10] &= 0xFF
        assert np.array_equal(sha(nw), g["bvh_nodes_sha"]), f"nodes differ (threads={threads})"
        assert np.array_equal(sha(idx), g["bvh_idx_sha"])


def test_bvh_builder_identical_live(ref):
    """a mid-size scene against the live reference builder, incl. duplicate centroids and degenerate tris"""
    from elevenrender_b200 import scenes
    tris = scenes.sphere_field(side=4, lon=20, lat=10, spacing=0.4, radius=0.15, z0=2.0)
    tris = np.concatenate([tris, tris[:50], scenes.displaced_grid(30, 20)], axis=0)
    rs = ref.RefScene(8, 8)
    rs.add_tris(tris)
    rs.finalize()
    rn, ri = rs.get_bvh()
    nodes, idx = capi.bvh_build(tris, 8)
    a = rn.view(np.uint32).reshape(-1, 11)
    b = nodes.view(np.uint32).reshape(-1, 11)
    assert np.array_equal(ri, idx)
    assert np.array_equal(a[:, :10], b[:, :10])
    assert np.array_equal(a[:, 10] & 0xFF, b[:, 10] & 0xFF)
    rs.close()


@pytest.mark.parametrize("name", ["grid", "textured", "cornell"])
def test_hdri_cdf_identical(name):
    g = golden(name)
    sc = common.to_ours(common.make_spec(name))
    assert np.array_equal(sha(sc.hdri.cdf), g["cdf_sha"])
    assert np.float32(sc.hdri.radianceSum) == g["radianceSum"]


@pytest.mark.parametrize("name", ["grid", "textured", "spheres", "cornell"])
def test_flatten_roundtrip(name):
    """T2: AoS -> device pair-record / SoA layout -> back is lossless"""
    spec = common.make_spec(name)
    sc = common.to_ours(spec)
    sc.buildBVH(4)
    nodes, slots, slack = host.flatten_roundtrip(sc)
    src = sc.nodes
    d = src["depth"]
    inner = d >= 1
    assert np.array_equal(nodes["b1"][inner].view(np.uint32), src["b1"][inner].view(np.uint32))
    assert np.array_equal(nodes["b2"][inner].view(np.uint32), src["b2"][inner].view(np.uint32))
    leaf = d == 18
    assert np.array_equal(nodes["from"][leaf], src["from"][leaf])
    assert np.array_equal(nodes["to"][leaf], src["to"][leaf])
    want = sc.tris[sc.triIndices]
    # uv.z is not carried (never read on the path: Tri.h:76 only feeds tUV.x/.y into Hit.tu/tv)
    keep = np.ones(39, dtype=bool)
    keep[[11, 14, 17]] = False
    assert np.array_equal(slots.view(np.uint32)[:, keep], want.view(np.uint32)[:, keep])
    assert np.isfinite(slack) and slack > 0


def _node_sha(nodes):
    nw = np.ascontiguousarray(nodes).view(np.uint32).reshape(-1, 11).copy()
    nw[:, 10] &= 0xFF
    return sha(nw)


def test_bvh_builder_identical_c2_live(ref):
    """BASELINE config 2's full mesh (249,924 triangles): er_bvh_build against the LIVE reference builder
    (BVH::build, ~6 s single-threaded) — the arrays themselves, and the committed checksum of that run."""
    from elevenrender_b200 import workloads
    tris = workloads.workload("c2")["tris"]
    rs = ref.RefScene(8, 8)
    rs.add_tris(tris)
    rs.finalize()
    rn, ri = rs.get_bvh()
    rs.close()
    nodes, idx = capi.bvh_build(tris, 0)
    a = rn.view(np.uint32).reshape(-1, 11)
    b = nodes.view(np.uint32).reshape(-1, 11)
    assert np.array_equal(ri, idx)
    assert np.array_equal(a[:, :10], b[:, :10])
    assert np.array_equal(a[:, 10] & 0xFF, b[:, 10] & 0xFF)
    g = np.load(os.path.join(common.GOLDEN_DIR, "bvh_big.npz"))
    assert np.array_equal(_node_sha(nodes), g["c2_nodes_sha"]) and np.array_equal(sha(idx), g["c2_idx_sha"])


def test_bvh_builder_identical_c3_checksum():
    """BASELINE config 3's full mesh (10,000,000 triangles): er_bvh_build against the checksum of the
    reference builder's own arrays (tests/golden/make_bvh_golden.py ran BVH::build once: 87 s, one thread)."""
    from elevenrender_b200 import workloads
    g = np.load(os.path.join(common.GOLDEN_DIR, "bvh_big.npz"))
    tris = workloads.workload("c3")["tris"]
    assert tris.shape[0] == int(g["c3_tris"]) == 10_000_000
    nodes, idx = capi.bvh_build(tris, 0)
    assert np.array_equal(_node_sha(nodes), g["c3_nodes_sha"]), "nodes[] differ from BVH::build at 10 M triangles"
    assert np.array_equal(sha(idx), g["c3_idx_sha"]), "triIndices[] differ from BVH::build at 10 M triangles"
