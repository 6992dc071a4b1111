mkdir -p gpurun_out
(timeout 300 python -m pytest tests/test_gpu_wavefront.py -q -x) 2>&1 | tail -2
{
for w in c2 c3; do
KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "watchdog report out of line|" 2>&1 | tail -1
ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_wdinline.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "watchdog report inline (before)|" 2>&1 | tail -1
done
} > gpurun_out/r04z_watchdog.txt 2>&1
cat gpurun_out/r04z_watchdog.txt
