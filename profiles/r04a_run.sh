# round 2, final evidence run 1 of the restored tree: whole GPU suite, the default bench line, the launch list
mkdir -p gpurun_out
rm -f gpurun_out/parity_outliers.jsonl
(timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/r04a_pytest.log 2>&1; rc=$?
tail -5 gpurun_out/r04a_pytest.log
cp gpurun_out/parity_outliers.jsonl gpurun_out/r04a_parity_outliers.jsonl 2>/dev/null
[ $rc -ne 0 ] && exit 1
timeout 600 python bench.py > gpurun_out/r04a_bench_n1.json 2> gpurun_out/r04a_bench_n1.err || { tail -5 gpurun_out/r04a_bench_n1.err; exit 1; }
cat gpurun_out/r04a_bench_n1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r04a_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r04a_ncu_launch.log 2>&1
tail -3 gpurun_out/r04a_launches.csv
