# final build: whole GPU suite, smoke(), the default bench line and the reference arm, launch list
mkdir -p gpurun_out
rm -f gpurun_out/parity_outliers.jsonl
(timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/r04x_pytest.log 2>&1; rc=$?
tail -4 gpurun_out/r04x_pytest.log
cp gpurun_out/parity_outliers.jsonl gpurun_out/r04x_parity_outliers.jsonl 2>/dev/null
[ $rc -ne 0 ] && exit 1
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/r04x_bench_n1.json 2> gpurun_out/r04x_bench_n1.err || { tail -5 gpurun_out/r04x_bench_n1.err; exit 1; }
python -c "
import json; d=json.loads(open('gpurun_out/r04x_bench_n1.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'],'cpu',d['cpu_baseline']['value'],'c2',d['secondary']['c2']['value'],'clocks',d['clocks'],'launches',d['gpu_launches'])"
timeout 600 python bench.py --impl reference > gpurun_out/r04x_bench_reference_n1.json 2>/dev/null; cut -c1-300 gpurun_out/r04x_bench_reference_n1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r04x_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r04x_ncu_launch.log 2>&1
tail -2 gpurun_out/r04x_launches.csv
