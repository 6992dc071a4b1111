# 4 GPUs: the bench line at N=4 (strong scaling, C3 at 256 spp) with the feedback balance
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29557 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r05f_bench_n4.json 2> gpurun_out/r05f_bench_n4.err || tail -20 gpurun_out/r05f_bench_n4.err
python -c "
import json; d=json.loads(open('gpurun_out/r05f_bench_n4.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],d['ms_per_step_fastest_rank'],'t6',d['t6'],'stripes',d['stripes']); print(d['rebalance']); print('c2',d['secondary']['c2']['value'])"
