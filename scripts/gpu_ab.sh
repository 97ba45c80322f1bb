#!/bin/bash
# Kernel ms of library variants (raytracing-project_b200/variants/<v>.so; "default" = the tree's build) on a few scenes.
# Usage under gpurun: bash scripts/gpu_ab.sh <tag> "<scene[:paper]> ..." "<variant[@opt=val,...]> ..."
TAG=$1; SCENES=$2; VARS=$3
OUT=gpurun_out/$TAG; mkdir -p $OUT
cp raytracing-project_b200/librt_b200.so /tmp/default.so
for vv in $VARS; do
  v=${vv%%@*}; opt=""; [[ "$vv" == *@* ]] && opt=${vv##*@}
  if [ "$v" = default ]; then cp /tmp/default.so raytracing-project_b200/librt_b200.so; else cp raytracing-project_b200/variants/$v.so raytracing-project_b200/librt_b200.so; fi
  for a in $SCENES; do
    sc=${a%%:*}; mode=standard; [[ "$a" == *:* ]] && mode=${a##*:}
    RT_OPT=$opt timeout -k 10 300 python scripts/profile_run.py scenes/$sc.json $mode 4 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read().strip()); print('$vv','$a','kernel_ms',round(d['kernel_ms'],3),'tests/ray',round(d['prim_tests']/(d['primary_rays']+d['shadow_rays']+d['secondary_rays']),2))" | tee -a $OUT/ab.txt
  done
done
cp /tmp/default.so raytracing-project_b200/librt_b200.so
