mkdir -p gpurun_out
for cfg in "default|" "plain|ER_WF_CHUNKED=0" "chunk4|ER_WF_CHUNK_SPP=4" "chunk32|ER_WF_CHUNK_SPP=32"; do
  label="${cfg%%|*}"; envs="${cfg#*|}"
  echo "== $label ($envs)"
  env $envs timeout 600 python profiles/shard_emulate.py c3 8 256 stripes 2>&1 | tail -3
done > gpurun_out/r02q_shards_c3.txt 2>&1
cat gpurun_out/r02q_shards_c3.txt
