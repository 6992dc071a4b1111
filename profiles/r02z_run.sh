mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02z_bench_short.json 2> gpurun_out/r02z_bench_short.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02z_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02z_ncu_launch.log 2>&1
KERNEL=0 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_render_wavefront --launch-count 1 -o gpurun_out/r02z_c2_full -f python profiles/quick.py c2 16 "f|" > gpurun_out/r02z_ncu_c2.log 2>&1
KERNEL=0 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:k_render_wavefront --launch-count 1 -o gpurun_out/r02z_c3_full -f python profiles/quick.py c3 8 "f|" > gpurun_out/r02z_ncu_c3.log 2>&1
ls -la gpurun_out/r02z*
