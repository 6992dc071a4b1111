"""Generates tests/golden/*.npz from the reference's own code (oracle/_ref/liber_ref.so, built from
/root/reference/src by oracle/Makefile).  Run in the build container (needs /root/reference to
build the oracle):   make -C oracle ref && python tests/golden/make_golden.py
The fixtures pin: the RNG streams, the BVH arrays of the reference builder (as checksums), HDRI CDFs,
nearest-hit records for host-generated rays, and fixed-seed passes / RNG state / sample counters.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import common  # noqa: E402
import refharness as R  # noqa: E402

SCENES = ["grid", "grid_rot", "cornell", "textured", "spheres", "miss"]
SPP = [1, 4]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    out = {}
    # RNG known answers (kernel.cpp:25-47)
    for idx in (0, 1, 12345, 2073599):
        st, v = R.rng_stream(idx, 16)
        out[f"rng_state_{idx}"] = st
        out[f"rng_value_{idx}"] = v
    out["sizeof"] = np.array(list(R.sizeof_table().values()), dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "kat.npz"), **out)

    for name in SCENES:
        spec = common.make_spec(name)
        rs = common.to_ref(spec)
        g = {}
        nodes, idx = rs.get_bvh()
        nw = nodes.view(np.uint32).reshape(-1, 11).copy()
        nw[:, 10] &= 0xFF  # bool valid + 3 padding bytes
        g["bvh_nodes_sha"] = np.frombuffer(bytes.fromhex(sha(nw)), dtype=np.uint8)
        g["bvh_idx_sha"] = np.frombuffer(bytes.fromhex(sha(idx)), dtype=np.uint8)
        hd = rs.get_hdri()
        g["cdf_sha"] = np.frombuffer(bytes.fromhex(sha(hd["cdf"])), dtype=np.uint8)
        g["radianceSum"] = np.float32(hd["radianceSum"])
        rays = common.primary_rays(spec)
        g["hits"] = rs.trace(rays).view(np.uint32)
        done = 0
        for spp in SPP:
            rs.render(spp - done, nthreads=0)
            done = spp
            p = rs.get_passes()
            g[f"passes_{spp}"] = p[[0, 2, 3, 4]]  # DENOISE is never written
            g[f"rng_{spp}"] = rs.get_rng()
            g[f"samples_{spp}"] = rs.get_samples()
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **g)
        print(name, "tris", spec["tris"].shape[0], "valid hits", int((g["hits"][:, 15] != 0).sum()), "/", rays.shape[0],
              "beauty mean", float(p[0, ..., :3].mean()))
        rs.close()


if __name__ == "__main__":
    main()
