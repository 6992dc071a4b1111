# plain vs chunked kernel over all stripes of the 4-way and 8-way split (C3, 256 spp), new scheduler defaults
mkdir -p gpurun_out
for n in 8 4; do
  echo "== N=$n chunked"; ER_WF_CHUNKED=1 timeout 900 python profiles/shard_emulate.py c3 $n 256 stripes 2>&1 | tail -1
  echo "== N=$n plain b256"; ER_WF_THR_BATCH=256 timeout 900 python profiles/shard_emulate.py c3 $n 256 stripes 2>&1 | tail -1
done > gpurun_out/r04l_shards.txt 2>&1
echo "== N=8 plain b384" >> gpurun_out/r04l_shards.txt; ER_WF_THR_BATCH=384 timeout 900 python profiles/shard_emulate.py c3 8 256 stripes 2>&1 | tail -1 >> gpurun_out/r04l_shards.txt
cat gpurun_out/r04l_shards.txt
