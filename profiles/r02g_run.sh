mkdir -p gpurun_out
KERNEL=3 REPS=1 ER_WF_THR_DESCEND=16 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_st_trace --launch-skip 30 --launch-count 1 -o gpurun_out/r02g_trace_c2 -f python profiles/quick.py c2 16 "k3|" > gpurun_out/r02g_ncu.log 2>&1
KERNEL=3 REPS=1 ER_WF_THR_DESCEND=16 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:k_st_ -c 1500 --csv --log-file gpurun_out/r02g_launches_c2.csv python profiles/quick.py c2 16 "k3|" > gpurun_out/r02g_ncu2.log 2>&1
