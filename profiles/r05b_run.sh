# final binaries (clean rebuild): whole GPU suite, smoke(), default bench line
mkdir -p gpurun_out
rm -f gpurun_out/parity_outliers.jsonl
(timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/r05b_pytest.log 2>&1; rc=$?
tail -3 gpurun_out/r05b_pytest.log
[ $rc -ne 0 ] && exit 1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/r05b_bench_n1.json 2> gpurun_out/r05b_bench_n1.err || { tail -5 gpurun_out/r05b_bench_n1.err; exit 1; }
python -c "
import json; d=json.loads(open('gpurun_out/r05b_bench_n1.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'],'cpu',d['cpu_baseline']['value'],'c2',d['secondary']['c2']['value'],d['secondary']['c2']['e2e']['value'],'clocks',d['clocks'])"
