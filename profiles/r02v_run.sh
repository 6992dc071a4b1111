mkdir -p gpurun_out
STRIPE=400,408 KERNEL=0 REPS=2 timeout 900 python profiles/quick.py c3 128 "base|ER_WF_CHUNKED=0" "b32|ER_WF_CHUNKED=0 ER_WF_THR_BATCH=32" "b64|ER_WF_CHUNKED=0 ER_WF_THR_BATCH=64" "sw4|ER_WF_CHUNKED=0 ER_WF_THR_SWITCH=4" "sw16|ER_WF_CHUNKED=0 ER_WF_THR_SWITCH=16" "sw32|ER_WF_CHUNKED=0 ER_WF_THR_SWITCH=32" "sw32b32|ER_WF_CHUNKED=0 ER_WF_THR_SWITCH=32 ER_WF_THR_BATCH=32" "d1|ER_WF_CHUNKED=0 ER_WF_THR_DESCEND=1" "d32|ER_WF_CHUNKED=0 ER_WF_THR_DESCEND=32" "coop64|ER_WF_CHUNKED=0 ER_WF_COOP_MIN=64" 2>&1 | tail -10 > gpurun_out/r02v_lowload.txt
STRIPE=400,408 KERNEL=0 REPS=2 timeout 900 python profiles/quick.py c2 128 "c2-base|ER_WF_CHUNKED=0" "c2-sw32|ER_WF_CHUNKED=0 ER_WF_THR_SWITCH=32" 2>&1 | tail -2 >> gpurun_out/r02v_lowload.txt
STRIPE=400,408 KERNEL=1 REPS=2 timeout 900 python profiles/quick.py c3 128 "c3-kernel1-pixel|" 2>&1 | tail -1 >> gpurun_out/r02v_lowload.txt
cat gpurun_out/r02v_lowload.txt
