# costly row blocks first (er_set_block_order): the 8-way and 4-way split of C3 at 256 spp, plain and chunked, with and without
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_wavefront.py tests/test_gpu_group.py -q -x) > gpurun_out/r04s_test.log 2>&1; tail -2 gpurun_out/r04s_test.log
{
for n in 8 4; do
  echo "== N=$n plain, block order"; ER_WF_CHUNKED=0 timeout 900 python profiles/shard_emulate.py c3 $n 256 stripes 2>&1 | tail -1
  echo "== N=$n plain, top to bottom"; BLOCK_ORDER=0 ER_WF_CHUNKED=0 timeout 900 python profiles/shard_emulate.py c3 $n 256 stripes 2>&1 | tail -1
  echo "== N=$n chunked, block order"; ER_WF_CHUNKED=1 timeout 900 python profiles/shard_emulate.py c3 $n 256 stripes 2>&1 | tail -1
done
echo "== N=8 chunked flat chunks, top to bottom"; BLOCK_ORDER=0 ER_WF_CHUNK_TAIL=1:1:0:1 ER_WF_CHUNKED=1 timeout 900 python profiles/shard_emulate.py c3 8 256 stripes 2>&1 | tail -1
echo "== N=2 plain, block order"; timeout 900 python profiles/shard_emulate.py c3 2 256 stripes 2>&1 | tail -1
} > gpurun_out/r04s_shards.txt 2>&1
cat gpurun_out/r04s_shards.txt
