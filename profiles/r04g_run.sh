# what a path waits for when the SM is nearly empty: ncu of the plain kernel with 32 and with 128 path slots per SM (C3, 2 spp)
mkdir -p gpurun_out
for n in 32 128; do
ER_WF_CHUNKED=0 ER_WF_SLOT_LIMIT=$n ER_WF_THR_BATCH=32 KERNEL=0 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_render_wavefront --launch-count 1 -o gpurun_out/r04g_c3_slots$n -f python profiles/quick.py c3 2 "f|" > gpurun_out/r04g_ncu_$n.log 2>&1
done
ls -la gpurun_out/r04g*
