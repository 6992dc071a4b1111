#!/bin/bash
# threshold sweep of k_render_wavefront on the bench workload (prints Msamples/s per setting)
for d in ${DS:-2 4 8 12 16}; do for s in ${SS:-1 4 8}; do
  v=$(ER_WF_WATCHDOG_MS=3000 ER_WF_THR_DESCEND=$d ER_WF_THR_SWITCH=$s timeout 90 python bench.py --steps 1 --warmup 3 --spp 16 --spp-per-launch 16 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1))")
  echo "thr_descend=$d thr_switch=$s Msamples/s=$v"
done; done
