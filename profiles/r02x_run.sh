mkdir -p gpurun_out
for w in c2 c3; do
  KERNEL=0 REPS=2 timeout 600 python profiles/quick.py $w 32 "s1024|" 2>&1 | tail -1
  for v in s768 s1280 s1536; do ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_$v.so KERNEL=0 REPS=2 timeout 600 python profiles/quick.py $w 32 "$v|" "$v-b512|ER_WF_THR_BATCH=512" 2>&1 | tail -2; done
done > gpurun_out/r02x_slots.txt 2>&1
cat gpurun_out/r02x_slots.txt
