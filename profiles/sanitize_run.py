#!/usr/bin/env python
"""small renders through every scheduling path of k_render_wavefront, meant to run under
`compute-sanitizer --tool memcheck` (slow: keep it tiny)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ["ER_WF_WATCHDOG_MS"] = "600000"
import numpy as np

import common
from elevenrender_b200 import host

def run(name, res, spp, env):
    for k, v in env.items():
        os.environ[k] = v
    spec = common.make_spec(name)
    spec["x_res"], spec["y_res"] = res
    rm = host.RenderingManager(host.RenderParameters(res[0], res[1], spp, spp_per_launch=spp))
    rm.upload(common.to_ours(spec))
    rm.render_samples(spp)
    img = rm.get_pass("beauty")
    rays = common.primary_rays(spec)[:2000]
    hits = rm.trace(rays)
    rm.close()
    for k in env:
        os.environ.pop(k, None)
    print(name, res, spp, env, "mean", float(img[..., :3].mean()), "hits", int(hits["valid"].sum()), flush=True)

run("textured", (96, 64), 3, {})
run("spheres", (96, 64), 6, {"ER_WF_THR_BATCH": "96"})
run("grid", (96, 64), 3, {"ER_WF_COOP_MIN": "1", "ER_WF_THR_BATCH": "32"})
run("grid_rot", (64, 48), 2, {"ER_WF_THR_DESCEND": "1", "ER_WF_THR_SWITCH": "31"})
print("done")
