mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_gpu_render.py tests/test_gpu_wavefront.py tests/test_gpu_trace.py -q -x -k "not bounds_check") > gpurun_out/r02y_test.log 2>&1; rc=$?; tail -3 gpurun_out/r02y_test.log
[ $rc -ne 0 ] && exit 1
for w in c2 c3; do
  KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 32 "top9|" 2>&1 | tail -1
  for v in top0 top7 top8 top10; do ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_$v.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 32 "$v|" 2>&1 | tail -1; done
done > gpurun_out/r02y_top.txt 2>&1
cat gpurun_out/r02y_top.txt
