#!/usr/bin/env python
"""ONE GPU renders, one after the other, every shard of an N-way sharded job (the per-GPU work of
`bench.py --gpus N`, strong scaling: the same frame and spp): per-shard kernel time, and what the N-GPU job
would take (the slowest shard) against the unsharded frame.
   shard_emulate.py <workload> <N> <spp> [rows|stripes ...]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elevenrender_b200 import capi, host, workloads  # noqa: E402

name, N, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
modes = sys.argv[4:] or ["rows", "stripes"]
ROWS = int(os.environ.get("ROWS", "8"))
spec = workloads.workload(name)
scene = workloads.scene_from_spec(spec)
scene.buildBVH()
pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR, spp_per_launch=spp)
rm = host.RenderingManager(pars)
rm.upload(scene)
P = spec["x_res"] * spec["y_res"]


def timed():
    best = 1e30
    for _ in range(2):
        rm.reset()
        rm.reset_counters()
        rm.render_samples(spp)
        ms, _ = rm.kernel_time()
        best = min(best, ms)
    return best


_forced = os.environ.pop("ER_WF_CHUNKED", None)  # the full frame: the library's own choice of kernel
full = timed()
if _forced is not None:
    os.environ["ER_WF_CHUNKED"] = _forced
print(f"{name} full frame {spp} spp: {full:.2f} ms = {P * spp / full / 1e3:.1f} Msamples/s", flush=True)
for mode in modes:
    if mode == "stripes":
        cost = rm.row_cost(0, int(os.environ.get("COST_STRIDE", "4")))
        yb = capi.balance_stripes(cost, ROWS, N)
        print("stripe boundaries", [int(v) for v in yb], flush=True)
    times = []
    for k in range(N):
        if mode == "rows":
            rm.set_shard(k, N, ROWS)
        else:
            rm.set_stripe(int(yb[k]), int(yb[k + 1]), ROWS)
            if os.environ.get("BLOCK_ORDER", "0") != "0" and yb[k + 1] > yb[k]:  # BLOCK_ORDER=1: costly row blocks first
                rm.set_block_order(capi.order_blocks(cost, ROWS, int(yb[k]), int(yb[k + 1])))
        times.append(timed())
    rm.set_shard(0, 1, ROWS)
    t = np.array(times)
    print(f"{name} N={N} {mode}: per-shard ms {np.round(t, 1).tolist()}  max {t.max():.2f}  sum {t.sum():.2f}  "
          f"speed-up over 1 GPU {full / t.max():.2f}x (balance {t.mean() / t.max():.2f}, work inflation {t.sum() / full:.2f}x)",
          flush=True)
rm.close()
