mkdir -p gpurun_out
for cfg in "c32b256|ER_WF_CHUNK_SPP=32 ER_WF_THR_BATCH=256" "c64b256|ER_WF_CHUNK_SPP=64 ER_WF_THR_BATCH=256" "c16b256|ER_WF_CHUNK_SPP=16 ER_WF_THR_BATCH=256" "c32b192|ER_WF_CHUNK_SPP=32 ER_WF_THR_BATCH=192"; do
  label="${cfg%%|*}"; envs="${cfg#*|}"
  echo "== $label ($envs)"
  env $envs timeout 600 python profiles/shard_emulate.py c3 8 256 stripes 2>&1 | tail -1
done > gpurun_out/r02w_shards_c3.txt 2>&1
echo "== c2 default" >> gpurun_out/r02w_shards_c3.txt
timeout 600 python profiles/shard_emulate.py c2 8 64 stripes 2>&1 | tail -1 >> gpurun_out/r02w_shards_c3.txt
echo "== c2 c16b256" >> gpurun_out/r02w_shards_c3.txt
ER_WF_CHUNK_SPP=16 ER_WF_THR_BATCH=256 timeout 600 python profiles/shard_emulate.py c2 8 64 stripes 2>&1 | tail -1 >> gpurun_out/r02w_shards_c3.txt
cat gpurun_out/r02w_shards_c3.txt
