#!/usr/bin/env python
"""Fast A/B of launch-time switches of k_render_wavefront: the scene is built and uploaded ONCE, then every
setting renders `spp` samples of the full frame and the kernel time (CUDA events inside the library) is read.
   quick.py <workload> <spp> "label|ENV=V ENV=V" ...        (settings are environment variables read at launch)
Prints one line per setting: Msamples/s.  Not a bench number (one launch, no L2 flush) — relative only."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elevenrender_b200 import host, workloads  # noqa: E402

name, spp = sys.argv[1], int(sys.argv[2])
reps = int(os.environ.get("REPS", "2"))
spec = workloads.workload(name)
scene = workloads.scene_from_spec(spec)
t0 = time.time()
scene.buildBVH()
print(f"{name}: scene ready, BVH {time.time() - t0:.1f} s", flush=True)
pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR, spp_per_launch=spp,
                             kernel=int(os.environ.get("KERNEL", "0")))
P = spec["x_res"] * spec["y_res"]
touched = set()
rm = None
for setting in sys.argv[3:] or ["default|"]:
    label, _, envs = setting.partition("|")
    for k in touched:
        os.environ.pop(k, None)
    reupload = False
    for kv in envs.split():
        k, _, v = kv.partition("=")
        os.environ[k] = v
        touched.add(k)
        reupload = reupload or k in ("ER_L2_PERSIST", "ER_SLACK")
    if rm is None or reupload:
        if rm is not None:
            rm.close()
        rm = host.RenderingManager(pars)
        rm.upload(scene)
    # STRIPE=y0,y1 (environment, or per setting): render only these rows (one shard of a strong-scaling job)
    if os.environ.get("STRIPE"):
        y0, y1 = (int(v) for v in os.environ["STRIPE"].split(","))
    else:
        y0, y1 = 0, spec["y_res"]
    rm.set_stripe(y0, y1, 8)
    P = spec["x_res"] * (y1 - y0)
    best = 1e30
    for _ in range(reps):
        rm.reset()
        rm.reset_counters()
        rm.render_samples(spp)
        ms, n = rm.kernel_time()
        best = min(best, ms)
    print(f"{label:40s} {P * spp / best / 1e3:9.1f} Msamples/s   {best:9.2f} ms  ({envs})", flush=True)
rm.close()
