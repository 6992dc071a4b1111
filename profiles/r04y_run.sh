mkdir -p gpurun_out
{
for w in c2 c3; do
KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "probes compiled out (default)|" 2>&1 | tail -1
ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_withtimeline.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "probes compiled in|" 2>&1 | tail -1
done
} > gpurun_out/r04y_probes.txt 2>&1
cat gpurun_out/r04y_probes.txt
