for s in "grid 4" "spheres 4"; do timeout 120 python profiles/dbg_staged.py $s 2>&1 | tail -4; done
echo SIMPLE
for s in "grid 4" "spheres 4"; do ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_simple.so timeout 120 python profiles/dbg_staged.py $s 2>&1 | tail -4; done
ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_simple.so ER_ST_THR_LEAF=32 ER_ST_THR_REFILL=32 timeout 120 python profiles/dbg_staged.py grid 1 2>&1 | tail -4
ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_simple.so ER_ST_THR_LEAF=1 ER_ST_THR_REFILL=32 timeout 120 python profiles/dbg_staged.py grid 1 2>&1 | tail -4
ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_simple.so ER_ST_THR_LEAF=32 ER_ST_THR_REFILL=1 timeout 120 python profiles/dbg_staged.py grid 1 2>&1 | tail -4
