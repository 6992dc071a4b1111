# 2 GPUs: GPU suite on a 2-GPU box (device groups on two real devices), then the bench line at N=2 with the feedback balance
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_group.py tests/test_gpu_wavefront.py tests/test_zz_no_alloc.py -q -x) > gpurun_out/r04m_test.log 2>&1; tail -3 gpurun_out/r04m_test.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r04m_bench_n2.json 2> gpurun_out/r04m_bench_n2.err || tail -20 gpurun_out/r04m_bench_n2.err
python -c "
import json; d=json.loads(open('gpurun_out/r04m_bench_n2.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'ms',d['ms_per_step'],d['ms_per_step_fastest_rank'],'t6',d['t6'],'stripes',d['stripes']); print(d['rebalance']); print('c2',d['secondary']['c2']['value'])"
