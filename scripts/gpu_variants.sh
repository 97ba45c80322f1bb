#!/bin/bash
# Time each variant library on a few scenes (kernel ms only). Usage: gpu_variants.sh <tag> "<scenes>" variant...
TAG=$1; SCENES=$2; shift 2
OUT=gpurun_out/$TAG; mkdir -p $OUT
cp raytracing-project_b200/librt_b200.so /tmp/default.so
for v in "$@"; do
  cp raytracing-project_b200/variants/$v.so raytracing-project_b200/librt_b200.so
  for a in $SCENES; do
    sc=${a%%:*}; mode=standard; [[ "$a" == *:* ]] && mode=${a##*:}
    python scripts/profile_run.py scenes/$sc.json $mode 4 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read().strip()); print('$v','$a','kernel_ms',round(d['kernel_ms'],3))" | tee -a $OUT/variants.txt
  done
done
cp /tmp/default.so raytracing-project_b200/librt_b200.so
