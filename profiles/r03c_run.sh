mkdir -p gpurun_out
(timeout 1400 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_wavefront.py tests/test_gpu_staged.py tests/test_gpu_configs.py tests/test_gpu_units.py -q -x) > gpurun_out/r03c_test.log 2>&1; rc=$?; tail -4 gpurun_out/r03c_test.log
[ $rc -ne 0 ] && exit 1
for w in c3 c2; do
  KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "groupboxes|" "groupboxes-coop4|ER_WF_COOP_MIN=4" "groupboxes-d8|ER_WF_THR_DESCEND=8" 2>&1 | tail -3
  ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_nogb.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "brute-force|" 2>&1 | tail -1
done > gpurun_out/r03c_groupboxes.txt 2>&1
cat gpurun_out/r03c_groupboxes.txt
