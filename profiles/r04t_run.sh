mkdir -p gpurun_out
nproc > gpurun_out/r04t_upload_timing.txt; free -g | head -2 >> gpurun_out/r04t_upload_timing.txt
timeout 600 python profiles/upload_timing.py c3 2>> gpurun_out/r04t_upload_timing.txt >/dev/null
tail -45 gpurun_out/r04t_upload_timing.txt
