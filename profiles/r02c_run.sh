mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_staged.py -q -x) > gpurun_out/r02c_staged_test.log 2>&1; echo "rc=$?" >> gpurun_out/r02c_staged_test.log
tail -15 gpurun_out/r02c_staged_test.log
for w in c2 c3; do
  for k in 0 3; do KERNEL=$k REPS=2 timeout 600 python profiles/quick.py $w 16 "k$k|" 2>&1 | tail -1; done
  for v in st3 st2; do ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_$v.so KERNEL=3 REPS=2 timeout 600 python profiles/quick.py $w 16 "k3-$v|" 2>&1 | tail -1; done
done > gpurun_out/r02c_quick.txt 2>&1
cat gpurun_out/r02c_quick.txt
