mkdir -p gpurun_out
for w in c2 c3; do
  KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "glibc-bits|" 2>&1 | tail -1
  ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_cudalibm.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "cuda-libm|" 2>&1 | tail -1
done > gpurun_out/r03a_libm_cost.txt 2>&1
cat gpurun_out/r03a_libm_cost.txt
