mkdir -p gpurun_out
for h in 8 40 80 120 160 320; do
  y1=$((400+h))
  STRIPE=400,$y1 KERNEL=0 REPS=2 timeout 900 python profiles/quick.py c3 128 "rows$h-auto|" "rows$h-plain|ER_WF_CHUNKED=0" "rows$h-chunked|ER_WF_CHUNKED=1" 2>&1 | tail -3
done > gpurun_out/r02u_latency.txt 2>&1
cat gpurun_out/r02u_latency.txt
