mkdir -p gpurun_out
rm -f gpurun_out/parity_outliers.jsonl
(timeout 900 python -m pytest tests -m gpu -q -x) > gpurun_out/r05e_pytest.log 2>&1; tail -2 gpurun_out/r05e_pytest.log
