# round 2, final evidence run 2: the no-alloc test alone in a fresh process, the reference arm, what the exact libm
# costs, the 8-way stripes on one GPU, one ncu --set full capture of the render kernel on C3 and on C2
mkdir -p gpurun_out
(timeout 400 python -m pytest tests/test_zz_no_alloc.py -q -x) > gpurun_out/r04b_noalloc_alone.log 2>&1; tail -2 gpurun_out/r04b_noalloc_alone.log
(timeout 400 python -m pytest tests/test_gpu_translated.py tests/test_gpu_trace.py -q -x) > gpurun_out/r04b_translated.log 2>&1; tail -15 gpurun_out/r04b_translated.log
timeout 700 python bench.py --impl reference > gpurun_out/r04b_bench_reference_n1.json 2> gpurun_out/r04b_bench_reference_n1.err; cat gpurun_out/r04b_bench_reference_n1.json
for w in c3 c2; do
  KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "glibc-bit libm (shipped)|" 2>&1 | tail -1
  ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_cudalibm.so KERNEL=0 REPS=3 timeout 600 python profiles/quick.py $w 64 "CUDA libm (-DER_CUDA_LIBM)|" 2>&1 | tail -1
done > gpurun_out/r04b_libm_cost.txt 2>&1
cat gpurun_out/r04b_libm_cost.txt
timeout 900 python profiles/shard_emulate.py c3 8 256 stripes > gpurun_out/r04b_shards_c3.txt 2>&1; tail -3 gpurun_out/r04b_shards_c3.txt
KERNEL=0 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_render_wavefront --launch-count 1 -o gpurun_out/r04b_c3_full -f python profiles/quick.py c3 8 "f|" > gpurun_out/r04b_ncu_c3.log 2>&1
KERNEL=0 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_render_wavefront --launch-count 1 -o gpurun_out/r04b_c2_full -f python profiles/quick.py c2 16 "f|" > gpurun_out/r04b_ncu_c2.log 2>&1
ls -la gpurun_out/r04b*
