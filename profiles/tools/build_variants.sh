#!/bin/bash
# Builds libgpt_b200.so variants HERE (nvcc cross-compiles) into variants/<name>.so so that one gpurun call can A/B them
# without spending box time on compiling.  usage: build_variants.sh name1 "flags1" name2 "flags2" ...
set -e
cd "$(dirname "$0")/../../g.p.u-pathtracer_b200"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  touch csrc/*.cu
  make EXTRA="$flags" > /tmp/build_$name.log 2>&1 || { tail -20 /tmp/build_$name.log; exit 1; }
  cp libgpt_b200.so ../variants/$name.so
  echo "$name: $flags"; grep -A2 "k_wf_shade\|k_wf_extend_wwILb1ELb0" csrc/gpt_wavefront.o.ptxas.log | grep -E "Used|spill" | head -4
done
touch csrc/*.cu
make > /dev/null 2>&1
