#!/bin/bash
# A/B of library variants / env settings on the bench workload: ab.sh "label|ENV=.. ENV=.." ...
for spec in "$@"; do
  label="${spec%%|*}"; envs="${spec#*|}"
  v=$(env $envs ER_WF_WATCHDOG_MS=3000 timeout 90 python bench.py --steps 2 --warmup 3 --spp 16 --spp-per-launch 16 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'Ms/s', round(d['mrays_per_s']), 'Mrays/s box/s', round(d['counters_per_sample']['box_tests'],1), 'tri/s', round(d['counters_per_sample']['tri_tests'],2))")
  echo "$label: $v"
done
