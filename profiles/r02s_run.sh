mkdir -p gpurun_out
STRIPE=128,272 KERNEL=0 REPS=2 timeout 900 python profiles/quick.py c3 128 "full-default|" "b256|ER_WF_THR_BATCH=256" "b128|ER_WF_THR_BATCH=128" "b64|ER_WF_THR_BATCH=64" "b32|ER_WF_THR_BATCH=32" "plain|ER_WF_CHUNKED=0" "plain-b128|ER_WF_CHUNKED=0 ER_WF_THR_BATCH=128" "plain-b64|ER_WF_CHUNKED=0 ER_WF_THR_BATCH=64" "d8|ER_WF_THR_DESCEND=8" "d24|ER_WF_THR_DESCEND=24" 2>&1 | tail -10 > gpurun_out/r02s_stripe_knobs.txt
KERNEL=0 REPS=2 timeout 900 python profiles/quick.py c3 32 "fullframe-default|" "fullframe-b128|ER_WF_THR_BATCH=128" 2>&1 | tail -2 >> gpurun_out/r02s_stripe_knobs.txt
STRIPE=128,272 KERNEL=3 REPS=2 timeout 900 python profiles/quick.py c3 128 "staged|" 2>&1 | tail -1 >> gpurun_out/r02s_stripe_knobs.txt
cat gpurun_out/r02s_stripe_knobs.txt
