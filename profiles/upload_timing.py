#!/usr/bin/env python
"""where er_upload_scene spends its time on a workload (ER_TIMING laps on stderr): upload_timing.py c3"""
import os
import sys
import time

os.environ["ER_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elevenrender_b200 import host, workloads  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
spec = workloads.workload(name)
scene = workloads.scene_from_spec(spec)
scene.buildBVH()
rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], 1))
for i in range(4):
    t0 = time.perf_counter()
    rm.upload(scene)
    rm.sync()
    print(f"upload {i}: {1e3 * (time.perf_counter() - t0):.1f} ms", file=sys.stderr, flush=True)
rm.close()
