#!/bin/bash
# Build experimental variants of librt_b200.so into raytracing-project_b200/variants/<name>.so
# Usage: bash scripts/build_variants.sh name1="-DFLAG=1 ..." name2="..."
set -e
cd "$(dirname "$0")/../raytracing-project_b200"
mkdir -p variants
for spec in "$@"; do
  name=${spec%%=*}; flags=${spec#*=}
  rm -f build/rt_api.o
  make -s EXTRA="$flags" librt_b200.so
  cp librt_b200.so variants/$name.so
  grep -A3 "render_standard_kernelILb1" build/ptxas.log | grep -E "Used|spill" | sed "s/^/[$name] /"
done
rm -f build/rt_api.o; make -s   # restore the default build
