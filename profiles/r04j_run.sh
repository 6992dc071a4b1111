# batch size with the adaptive threshold and parking: full frame and stripes (C3, 256 spp), plain kernel
mkdir -p gpurun_out
B="ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=64 ER_WF_PARK_LANES=24"
KERNEL=0 REPS=2 timeout 1400 python profiles/quick.py c3 256 \
  "full b384|$B" "full b320|$B ER_WF_THR_BATCH=320" "full b256|$B ER_WF_THR_BATCH=256" \
  "1/2 b384|$B STRIPE=0,544" "1/2 b256|$B ER_WF_THR_BATCH=256 STRIPE=0,544" \
  "1/4 b384|$B STRIPE=272,544" "1/4 b256|$B ER_WF_THR_BATCH=256 STRIPE=272,544" \
  "1/8 b384|$B STRIPE=272,416" "1/8 b320|$B ER_WF_THR_BATCH=320 STRIPE=272,416" "1/8 b256|$B ER_WF_THR_BATCH=256 STRIPE=272,416" "1/8 b192|$B ER_WF_THR_BATCH=192 STRIPE=272,416" \
  "1/8 b256 bmin48|$B ER_WF_THR_BATCH=256 ER_WF_BATCH_MIN=48 STRIPE=272,416" "1/8 b256 bmin96|$B ER_WF_THR_BATCH=256 ER_WF_BATCH_MIN=96 STRIPE=272,416" \
  "1/8 b256 park20|$B ER_WF_THR_BATCH=256 ER_WF_PARK_LANES=20 STRIPE=272,416" "1/8 b256 park28|$B ER_WF_THR_BATCH=256 ER_WF_PARK_LANES=28 STRIPE=272,416" \
  "1/8 b256 reserve704|$B ER_WF_THR_BATCH=256 ER_WF_BATCH_RESERVE=704 STRIPE=272,416" "1/8 b256 reserve768|$B ER_WF_THR_BATCH=256 ER_WF_BATCH_RESERVE=768 STRIPE=272,416" \
  > gpurun_out/r04j_stripes.txt 2>&1
cat gpurun_out/r04j_stripes.txt
KERNEL=0 REPS=3 timeout 600 python profiles/quick.py c2 64 "c2 full base|ER_WF_CHUNKED=0 ER_WF_BATCH_MIN=384" "c2 full b384 bmin64 park24|$B" "c2 full b256|$B ER_WF_THR_BATCH=256" \
   "c2 1/8 chunked|ER_WF_CHUNKED=1 ER_WF_BATCH_MIN=256 STRIPE=272,416" "c2 1/8 plain b384|$B STRIPE=272,416" "c2 1/8 plain b256|$B ER_WF_THR_BATCH=256 STRIPE=272,416" > gpurun_out/r04j_c2.txt 2>&1
cat gpurun_out/r04j_c2.txt
