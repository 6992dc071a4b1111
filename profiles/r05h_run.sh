# chunked kernel on the slowest 1/8 stripe of C3 (rows 120-272, 256 spp): parking / batch knobs now that waiting slots are not live
mkdir -p gpurun_out
B="ER_WF_CHUNKED=1 STRIPE=120,272"
KERNEL=0 REPS=2 timeout 300 python profiles/quick.py c3 256 "default (park24 bmin64 b256)|$B" "park16|$B ER_WF_PARK_LANES=16" "park32|$B ER_WF_PARK_LANES=32" \
  "bmin32|$B ER_WF_BATCH_MIN=32" "bmin128|$B ER_WF_BATCH_MIN=128" "b192|$B ER_WF_THR_BATCH=192" "b320|$B ER_WF_THR_BATCH=320" "park0|$B ER_WF_PARK_LANES=0" \
  > gpurun_out/r05h_chunked_knobs.txt 2>&1
cat gpurun_out/r05h_chunked_knobs.txt
