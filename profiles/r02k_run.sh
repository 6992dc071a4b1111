mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_gpu_render.py tests/test_gpu_wavefront.py tests/test_gpu_staged.py tests/test_gpu_trace.py -q -x) > gpurun_out/r02k_test.log 2>&1; rc=$?; echo "rc=$rc" >> gpurun_out/r02k_test.log
tail -5 gpurun_out/r02k_test.log
[ $rc -ne 0 ] && exit 1
for w in c2 c3; do
  ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_nolean.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 32 "k0-nolean|" 2>&1 | tail -1
  KERNEL=0 REPS=3 timeout 900 python profiles/quick.py $w 32 "k0-lean|" "k0-lean-d12|ER_WF_THR_DESCEND=12" "k0-lean-d16|ER_WF_THR_DESCEND=16" "k0-lean-d4|ER_WF_THR_DESCEND=4" 2>&1 | tail -4
done > gpurun_out/r02k_quick.txt 2>&1
cat gpurun_out/r02k_quick.txt
