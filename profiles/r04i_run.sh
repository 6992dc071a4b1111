# adaptive batch threshold / parking on the full frame and on 1/2, 1/4 stripes (C3, 256 spp): which kernel for which shard size
mkdir -p gpurun_out
KERNEL=0 REPS=2 timeout 1400 python profiles/quick.py c3 256 \
  "full plain base|ER_WF_CHUNKED=0" \
  "full plain bmin64 park24|ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=64 ER_WF_PARK_LANES=24" \
  "full plain bmin64 park16|ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=64 ER_WF_PARK_LANES=16" \
  "1/2 (0-544) plain base|ER_WF_CHUNKED=0 STRIPE=0,544" \
  "1/2 (0-544) plain bmin64 park24|ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=64 ER_WF_PARK_LANES=24 STRIPE=0,544" \
  "1/2 (0-544) chunked|ER_WF_CHUNKED=1 STRIPE=0,544" \
  "1/4 (272-544) plain bmin64 park24|ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=64 ER_WF_PARK_LANES=24 STRIPE=272,544" \
  "1/4 (272-544) chunked|ER_WF_CHUNKED=1 STRIPE=272,544" \
  "1/4 (272-544) chunked bmin64 park24|ER_WF_CHUNKED=1 ER_WF_BATCH_MIN=64 ER_WF_PARK_LANES=24 STRIPE=272,544" \
  "1/8 plain bmin32 park24|ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=32 ER_WF_PARK_LANES=24 STRIPE=272,416" \
  "1/8 plain bmin128 park24|ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=128 ER_WF_PARK_LANES=24 STRIPE=272,416" \
  "1/8 plain bmin64 park24 reserve512|ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=64 ER_WF_BATCH_RESERVE=512 ER_WF_PARK_LANES=24 STRIPE=272,416" \
  "1/8 plain b256 bmin64 park24|ER_WF_CHUNKED=0 ER_WF_THR_BATCH=256 ER_WF_BATCH_MIN=64 ER_WF_PARK_LANES=24 STRIPE=272,416" \
  "1/8 plain b128 bmin64 park24|ER_WF_CHUNKED=0 ER_WF_THR_BATCH=128 ER_WF_BATCH_MIN=64 ER_WF_PARK_LANES=24 STRIPE=272,416" \
  > gpurun_out/r04i_stripes.txt 2>&1
cat gpurun_out/r04i_stripes.txt
