#!/usr/bin/env python
"""per-sample work counters of a workload (count_stats build of the same kernel): count.py <workload> <spp>"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elevenrender_b200 import host, workloads
name, spp = sys.argv[1], int(sys.argv[2])
spec = workloads.workload(name)
scene = workloads.scene_from_spec(spec)
scene.buildBVH()
pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR, spp_per_launch=spp, count_stats=True,
                             kernel=int(os.environ.get("KERNEL", "0")))
rm = host.RenderingManager(pars)
rm.upload(scene)
rm.reset_counters()
rm.render_samples(spp)
c = rm.get_counters()
print(name, {k: round(v / max(1, c["samples"]), 2) for k, v in c.items()})
rm.close()
