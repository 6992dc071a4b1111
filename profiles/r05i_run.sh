mkdir -p gpurun_out
(timeout 100 python -m pytest tests -m gpu -q -x) > gpurun_out/r05i_pytest.log 2>&1; tail -2 gpurun_out/r05i_pytest.log
