# chunked kernel: slots waiting for a chunk do not count as live paths (adaptive batch threshold / parking in the drain)
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_wavefront.py -q -x) 2>&1 | tail -1
{
echo "== N=8 chunked, live excludes waiting slots"; ER_WF_CHUNKED=1 timeout 900 python profiles/shard_emulate.py c3 8 256 stripes 2>&1 | tail -1
echo "== N=8 chunked, before"; ER_CUDA_LIB=$PWD/elevenrender_b200/variants/liber_livefull.so ER_WF_CHUNKED=1 timeout 900 python profiles/shard_emulate.py c3 8 256 stripes 2>&1 | tail -1
} > gpurun_out/r05d_chunked_live.txt 2>&1
cat gpurun_out/r05d_chunked_live.txt
