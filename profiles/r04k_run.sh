# adaptive batch threshold + warp parking as defaults, plain kernel at every shard size: whole GPU suite, bench line, stripes
mkdir -p gpurun_out
rm -f gpurun_out/parity_outliers.jsonl
(timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/r04k_pytest.log 2>&1; rc=$?
tail -5 gpurun_out/r04k_pytest.log
cp gpurun_out/parity_outliers.jsonl gpurun_out/r04k_parity_outliers.jsonl 2>/dev/null
[ $rc -ne 0 ] && exit 1
timeout 600 python bench.py > gpurun_out/r04k_bench_n1.json 2> gpurun_out/r04k_bench_n1.err || { tail -5 gpurun_out/r04k_bench_n1.err; exit 1; }
python -c "
import json; d=json.loads(open('gpurun_out/r04k_bench_n1.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'],'c2',d['secondary']['c2']['value'],'clocks',d['clocks'])"
timeout 900 python profiles/shard_emulate.py c3 8 256 stripes > gpurun_out/r04k_shards_c3.txt 2>&1; tail -3 gpurun_out/r04k_shards_c3.txt
timeout 900 python profiles/shard_emulate.py c3 4 256 stripes > gpurun_out/r04k_shards4_c3.txt 2>&1; tail -1 gpurun_out/r04k_shards4_c3.txt
timeout 900 python profiles/shard_emulate.py c3 2 256 stripes > gpurun_out/r04k_shards2_c3.txt 2>&1; tail -1 gpurun_out/r04k_shards2_c3.txt
