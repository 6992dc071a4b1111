# needs a library built with the probes: make -C elevenrender_b200/csrc EXTRA=-DWF_TIMELINE  (WF_TIMELINE build)
# which pixels finish last on a 1/8 stripe (plain kernel, so a pixel runs from its start to its end in one slot)
mkdir -p gpurun_out
ER_WF_TIMELINE=$PWD/gpurun_out/r04r_tl.bin ER_WF_CHUNKED=0 STRIPE=272,416 KERNEL=0 REPS=1 timeout 600 python profiles/quick.py c3 256 "tl|" > /dev/null 2>&1
python profiles/pixel_done.py gpurun_out/r04r_tl.bin > gpurun_out/r04r_pixel_done.txt; cat gpurun_out/r04r_pixel_done.txt; rm -f gpurun_out/r04r_tl.bin
