# where do the 1/8 stripes lose their 1.41x? plain vs chunked kernel on the full frame, on a 1/8 and on a 1/4 stripe (C3, 256 spp)
mkdir -p gpurun_out
KERNEL=0 REPS=2 timeout 1200 python profiles/quick.py c3 256 \
  "full plain|ER_WF_CHUNKED=0" "full chunked|ER_WF_CHUNKED=1" \
  "1/8 (272-416) plain|ER_WF_CHUNKED=0 STRIPE=272,416" "1/8 (272-416) chunked|ER_WF_CHUNKED=1 STRIPE=272,416" \
  "1/4 (272-544) plain|ER_WF_CHUNKED=0 STRIPE=272,544" "1/4 (272-544) chunked|ER_WF_CHUNKED=1 STRIPE=272,544" \
  "1/16 (272-344) plain|ER_WF_CHUNKED=0 STRIPE=272,344" "1/16 (272-344) chunked|ER_WF_CHUNKED=1 STRIPE=272,344" \
  "1/8 chunked, 64 spp chunks|ER_WF_CHUNKED=1 ER_WF_CHUNK_SPP=64 STRIPE=272,416" "1/8 chunked, 8 spp chunks|ER_WF_CHUNKED=1 ER_WF_CHUNK_SPP=8 STRIPE=272,416" \
  > gpurun_out/r04e_stripes.txt 2>&1
cat gpurun_out/r04e_stripes.txt
