mkdir -p gpurun_out
STRIPE=128,272 KERNEL=0 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_render_wavefront --launch-count 1 -o gpurun_out/r02t_c3_stripe -f python profiles/quick.py c3 32 "s|" > gpurun_out/r02t_ncu1.log 2>&1
KERNEL=0 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_render_wavefront --launch-count 1 -o gpurun_out/r02t_c3_full -f python profiles/quick.py c3 8 "f|" > gpurun_out/r02t_ncu2.log 2>&1
ls -la gpurun_out/r02t*.ncu-rep
