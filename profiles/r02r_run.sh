mkdir -p gpurun_out
for spp in 64 1024; do
  echo "== spp $spp"
  timeout 900 python profiles/shard_emulate.py c3 8 $spp stripes 2>&1 | tail -3
done > gpurun_out/r02r_shards_c3_spp.txt 2>&1
cat gpurun_out/r02r_shards_c3_spp.txt
