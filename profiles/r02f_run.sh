mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_staged.py -q -x) > gpurun_out/r02f_staged_test.log 2>&1; echo "rc=$?" >> gpurun_out/r02f_staged_test.log
tail -5 gpurun_out/r02f_staged_test.log
for w in c2 c3; do
  KERNEL=0 REPS=2 timeout 600 python profiles/quick.py $w 16 "k0|" 2>&1 | tail -1
  for v in v64 v80 v72 v96; do ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_$v.so KERNEL=3 REPS=2 timeout 600 python profiles/quick.py $w 16 "k3-$v-d8|ER_WF_THR_DESCEND=8" "k3-$v-d16|ER_WF_THR_DESCEND=16" "k3-$v-d24|ER_WF_THR_DESCEND=24" 2>&1 | tail -3; done
done > gpurun_out/r02f_quick.txt 2>&1
cat gpurun_out/r02f_quick.txt
