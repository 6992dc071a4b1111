mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_staged.py -q -x) > gpurun_out/r02h_staged_test.log 2>&1; rc=$?; echo "rc=$rc" >> gpurun_out/r02h_staged_test.log
tail -5 gpurun_out/r02h_staged_test.log
[ $rc -ne 0 ] && exit 1
for w in c2 c3; do
  KERNEL=0 REPS=2 timeout 600 python profiles/quick.py $w 16 "k0|" 2>&1 | tail -1
  KERNEL=3 REPS=2 timeout 900 python profiles/quick.py $w 16 "l12r12|" "l8r8|ER_ST_THR_LEAF=8 ER_ST_THR_REFILL=8" "l16r8|ER_ST_THR_LEAF=16 ER_ST_THR_REFILL=8" "l8r16|ER_ST_THR_LEAF=8 ER_ST_THR_REFILL=16" "l16r16|ER_ST_THR_LEAF=16 ER_ST_THR_REFILL=16" "l20r12|ER_ST_THR_LEAF=20 ER_ST_THR_REFILL=12" "l12r20|ER_ST_THR_LEAF=12 ER_ST_THR_REFILL=20" "l4r4|ER_ST_THR_LEAF=4 ER_ST_THR_REFILL=4" "l24r24|ER_ST_THR_LEAF=24 ER_ST_THR_REFILL=24" 2>&1 | tail -9
done > gpurun_out/r02h_quick.txt 2>&1
cat gpurun_out/r02h_quick.txt
