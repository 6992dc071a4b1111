mkdir -p gpurun_out
(timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -x) 2>&1 | tail -1
(timeout 600 python -m pytest tests/test_gpu_trace.py tests/test_gpu_group.py tests/test_gpu_upload.py tests/test_zz_no_alloc.py -q -x) 2>&1 | tail -1
