# software-pipelined cooperative leaf test (default build) against -DWF_NO_COOP_PIPE
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_wavefront.py tests/test_gpu_staged.py tests/test_gpu_configs.py -q -x) > gpurun_out/r04c_test.log 2>&1; rc=$?; tail -3 gpurun_out/r04c_test.log
[ $rc -ne 0 ] && exit 1
for w in c3 c2; do
  KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "coop pipelined (default)|" 2>&1 | tail -1
  ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_nopipe.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "no pipeline|" 2>&1 | tail -1
done > gpurun_out/r04c_pipe.txt 2>&1
cat gpurun_out/r04c_pipe.txt
