# fewer path slots per SM with the adaptive batch threshold (more L1): builds with WF_SLOTS=768 / 896 against 1024
mkdir -p gpurun_out
{
for w in c2 c3; do
KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "1024 slots (default)|" 2>&1 | tail -1
ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_s896.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "896 slots|" "896 slots, batch 192|ER_WF_THR_BATCH=192" 2>&1 | tail -2
ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_s768.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "768 slots|" "768 slots, batch 96|ER_WF_THR_BATCH=96" 2>&1 | tail -2
done
} > gpurun_out/r05a_slots.txt 2>&1
cat gpurun_out/r05a_slots.txt
