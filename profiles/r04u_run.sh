# 8 GPUs: where the end-to-end step goes (C2 only, short): ER_TIMING laps of rank 0 and the host side of the upload call
mkdir -p gpurun_out
nproc > gpurun_out/r04u_n8_timing.txt
ER_TIMING=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --workload c2 --steps 3 --warmup 3 > gpurun_out/r04u_bench_c2_n8.json 2> gpurun_out/r04u_bench_c2_n8.err
grep -E "er_upload_scene|flatten" gpurun_out/r04u_bench_c2_n8.err | tail -40 >> gpurun_out/r04u_n8_timing.txt
python -c "
import json; d=json.loads(open('gpurun_out/r04u_bench_c2_n8.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e'])" >> gpurun_out/r04u_n8_timing.txt
tail -30 gpurun_out/r04u_n8_timing.txt
