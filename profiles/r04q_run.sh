# shrinking chunks: tests of the chunked kernel, then the 1/8 stripe and the 8-way / 4-way split (C3, 256 spp)
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_wavefront.py tests/test_gpu_render.py -q -x) > gpurun_out/r04q_test.log 2>&1; rc=$?; tail -3 gpurun_out/r04q_test.log
[ $rc -ne 0 ] && exit 1
KERNEL=0 REPS=2 timeout 1200 python profiles/quick.py c3 256 \
  "1/8 chunked flat|ER_WF_CHUNKED=1 ER_WF_CHUNK_TAIL=1:1:0:1 STRIPE=272,416" \
  "1/8 chunked default tail .625:4:.25:8|ER_WF_CHUNKED=1 STRIPE=272,416" \
  "1/8 chunked .75:4:.125:8|ER_WF_CHUNKED=1 ER_WF_CHUNK_TAIL=0.75:4:0.125:8 STRIPE=272,416" \
  "1/8 chunked .75:4:.125:16|ER_WF_CHUNKED=1 ER_WF_CHUNK_TAIL=0.75:4:0.125:16 STRIPE=272,416" \
  "1/8 chunked .5:4:.25:16|ER_WF_CHUNKED=1 ER_WF_CHUNK_TAIL=0.5:4:0.25:16 STRIPE=272,416" \
  "1/8 chunked .5:2:.25:8|ER_WF_CHUNKED=1 ER_WF_CHUNK_TAIL=0.5:2:0.25:8 STRIPE=272,416" \
  "1/8 chunked k64 .5:4:.25:16|ER_WF_CHUNKED=1 ER_WF_CHUNK_SPP=64 ER_WF_CHUNK_TAIL=0.5:4:0.25:16 STRIPE=272,416" \
  "1/8 chunked .875:8:.0625:32|ER_WF_CHUNKED=1 ER_WF_CHUNK_TAIL=0.875:8:0.0625:32 STRIPE=272,416" \
  > gpurun_out/r04q_stripes.txt 2>&1
cat gpurun_out/r04q_stripes.txt
ER_WF_TIMELINE=$PWD/gpurun_out/r04q_tl.bin ER_WF_CHUNKED=1 STRIPE=272,416 KERNEL=0 REPS=1 timeout 600 python profiles/quick.py c3 256 "tl|" > /dev/null 2>&1
python profiles/timeline.py gpurun_out/r04q_tl.bin > gpurun_out/r04q_timeline.txt; tail -22 gpurun_out/r04q_timeline.txt; rm -f gpurun_out/r04q_tl.bin
