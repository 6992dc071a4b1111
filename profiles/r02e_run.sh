mkdir -p gpurun_out
KERNEL=3 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_st_trace --launch-skip 30 --launch-count 1 -o gpurun_out/r02e_trace_c2 -f python profiles/quick.py c2 16 "k3|" > gpurun_out/r02e_ncu.log 2>&1
KERNEL=3 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_st_shade --launch-skip 30 --launch-count 1 -o gpurun_out/r02e_shade_c2 -f python profiles/quick.py c2 16 "k3|" >> gpurun_out/r02e_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
