# latency vs occupancy: the full frame of C3 at 16 spp (plain kernel) with n path slots per SM in use
# per-sample latency = n * 148 / throughput
mkdir -p gpurun_out
KERNEL=0 REPS=2 timeout 1200 python profiles/quick.py c3 16 \
  "1024 slots|ER_WF_CHUNKED=0" \
  "768 slots, batch 288|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=768 ER_WF_THR_BATCH=288" \
  "768 slots, batch 128|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=768 ER_WF_THR_BATCH=128" \
  "512 slots, batch 192|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=512 ER_WF_THR_BATCH=192" \
  "512 slots, batch 64|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=512 ER_WF_THR_BATCH=64" \
  "256 slots, batch 96|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=256 ER_WF_THR_BATCH=96" \
  "256 slots, batch 32|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=256 ER_WF_THR_BATCH=32" \
  "128 slots, batch 32|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=128 ER_WF_THR_BATCH=32" \
  "64 slots, batch 32|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=64 ER_WF_THR_BATCH=32" \
  "32 slots, batch 32|ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=32 ER_WF_THR_BATCH=32" \
  > gpurun_out/r04f_occupancy.txt 2>&1
cat gpurun_out/r04f_occupancy.txt
