# needs a library built with the probes: make -C elevenrender_b200/csrc EXTRA=-DWF_TIMELINE  (WF_TIMELINE build)
# timeline of a 1/8 stripe of C3 at 256 spp: chunked and plain kernel (live paths, queues, tickets over time)
mkdir -p gpurun_out
for m in 1 0; do
ER_WF_TIMELINE=$PWD/gpurun_out/r04p_timeline_chunked$m.bin ER_WF_CHUNKED=$m STRIPE=272,416 KERNEL=0 REPS=1 timeout 600 python profiles/quick.py c3 256 "stripe chunked=$m|" 2>&1 | tail -1
python profiles/timeline.py gpurun_out/r04p_timeline_chunked$m.bin > gpurun_out/r04p_timeline_chunked$m.txt; cat gpurun_out/r04p_timeline_chunked$m.txt
done
rm -f gpurun_out/r04p_timeline_chunked*.bin
