#!/usr/bin/env python
"""one GPU renders ONE shard of a K-way sharded C2 job (the per-GPU work of `bench.py --gpus K`):
usage: shard_emulate.py K [spp]   -> kernel ms, per-GPU Msamples/s"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elevenrender_b200 import host, sharding, workloads

K = int(sys.argv[1])
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 64 * K  # total spp of the job
spec = workloads.workload("c2")
scene = workloads.scene_from_spec(spec)
scene.buildBVH()
pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR, spp_per_launch=spp)
ROWS = int(os.environ.get("ROWS", "8"))
INDEX = int(os.environ.get("INDEX", "0"))
rm = host.RenderingManager(pars, shard=(INDEX, K, ROWS))
rm.upload(scene)
own = sharding.owned_pixel_count(INDEX, K, spec["x_res"], spec["y_res"], ROWS)
for rep in range(2):
    rm.reset()
    rm.reset_counters()
    rm.render_samples(spp)
    ms, n = rm.kernel_time()
    print(f"K={K} idx={INDEX} rows={ROWS} spp={spp} owned={own} launches={n} kernel {ms:.2f} ms  {own * spp / ms / 1e3:.1f} Msamples/s per GPU", flush=True)
rm.close()
