"""Generates tests/golden/bvh_big.npz: SHA-256 of the reference builder's own nodes[] / triIndices[]
(BVH::build, src/BVH.cpp:132-281, run through oracle/_ref/liber_ref.so) on the full-size BASELINE meshes:
C2 (249,924 triangles, seconds) and C3 (10,000,000 triangles: minutes and several GB single-threaded, which
is why it is run once here and pinned as a checksum).  tests/test_oracle_cpu.py::test_bvh_builder_identical_c2
and ..._c3 compare er_bvh_build against these checksums.
   make -C oracle ref && python tests/golden/make_bvh_golden.py [c2] [c3]
"""
import hashlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import common  # noqa: E402,F401
import refharness as R  # noqa: E402
from elevenrender_b200 import workloads  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def node_words(nodes):
    nw = np.ascontiguousarray(nodes).view(np.uint32).reshape(-1, 11).copy()
    nw[:, 10] &= 0xFF  # bool valid + 3 padding bytes
    return nw


def main():
    names = sys.argv[1:] or ["c2", "c3"]
    path = os.path.join(HERE, "bvh_big.npz")
    out = dict(np.load(path)) if os.path.exists(path) else {}
    for name in names:
        spec = workloads.workload(name)
        rs = R.RefScene(8, 8)
        rs.add_tris(spec["tris"])
        t0 = time.time()
        rs.finalize()  # the reference's own BVH::build
        dt = time.time() - t0
        nodes, idx = rs.get_bvh()
        out[f"{name}_nodes_sha"] = np.frombuffer(bytes.fromhex(sha(node_words(nodes))), dtype=np.uint8)
        out[f"{name}_idx_sha"] = np.frombuffer(bytes.fromhex(sha(idx)), dtype=np.uint8)
        out[f"{name}_tris"] = np.int64(spec["tris"].shape[0])
        out[f"{name}_ref_build_s"] = np.float64(dt)
        print(name, spec["tris"].shape[0], "tris, reference BVH::build", round(dt, 1), "s", flush=True)
        rs.close()
        np.savez_compressed(path, **out)


if __name__ == "__main__":
    main()
