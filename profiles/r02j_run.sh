mkdir -p gpurun_out
for k in 0 3; do timeout 600 python bench.py --workload c2 --kernel $k --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02j_bench_c2_k$k.json 2> gpurun_out/r02j_bench_c2_k$k.err; python -c "
import json; d=json.load(open('gpurun_out/r02j_bench_c2_k$k.json')); print('k$k', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {a:round(b,2) for a,b in d['counters_per_sample'].items()})"; done
