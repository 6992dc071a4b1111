#!/bin/bash
# builds kernel variants for A/B runs: profiles/build_variants.sh name1:"flags" name2:"flags" ...
# -> elevenrender_b200/variants/liber_<name>.so (select with ER_CUDA_LIB); restores the default build last
set -e
cd "$(dirname "$0")/.."
mkdir -p elevenrender_b200/variants
for v in "$@"; do
  name="${v%%:*}"; flags="${v#*:}"
  touch elevenrender_b200/csrc/er_api.cu
  make -s -C elevenrender_b200/csrc EXTRA="$flags" >/dev/null
  cp elevenrender_b200/liber_cuda.so elevenrender_b200/variants/liber_$name.so
  echo "built $name ($flags): $(grep -A3 wavefrontILb0 elevenrender_b200/csrc/_build/ptxas.log | grep -o 'Used [0-9]* registers')"
done
touch elevenrender_b200/csrc/er_api.cu
make -s -C elevenrender_b200/csrc >/dev/null
