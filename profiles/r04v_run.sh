# device-side flatten for page-locked inputs: tests, upload timing, bench line
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_upload.py tests/test_gpu_render.py tests/test_gpu_group.py -q -x) > gpurun_out/r04v_test.log 2>&1; rc=$?; tail -15 gpurun_out/r04v_test.log
[ $rc -ne 0 ] && exit 1
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r04v_bench_n1.json 2> gpurun_out/r04v_bench_n1.err || { tail -5 gpurun_out/r04v_bench_n1.err; exit 1; }
python -c "
import json; d=json.loads(open('gpurun_out/r04v_bench_n1.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']); print('c2',d['secondary']['c2']['value'], d['secondary']['c2']['e2e'])"
