# ncu --set full of the render kernel, final build, C3 at 8 spp (refreshes profiles/ncu_traffic.json)
mkdir -p gpurun_out
KERNEL=0 REPS=1 timeout 400 ncu --set full --clock-control none --import-source on --kernel-name regex:k_render_wavefront --launch-count 1 -o gpurun_out/r05g_c3_full -f python profiles/quick.py c3 8 "f|" > gpurun_out/r05g_ncu_c3.log 2>&1
ls -la gpurun_out/r05g*
