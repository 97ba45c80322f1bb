#!/bin/bash
# ncu --set full capture of the trace kernel on given scenes (each after a plain run exited 0).
# Usage: bash scripts/gpu_prof.sh <tag> scene[:mode] ...
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
for a in "$@"; do
  sc=${a%%:*}; mode=standard; [[ "$a" == *:* ]] && mode=${a##*:}
  pat="render_(standard|wavefront)"; [[ $mode == paper ]] && pat=paper_gbuffer
  python scripts/profile_run.py scenes/$sc.json $mode 2 > $OUT/plain_$sc.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$pat" -s 1 -c 1 -o $OUT/prof_${sc}_$mode \
      python scripts/profile_run.py scenes/$sc.json $mode 2 > $OUT/ncu_$sc.log 2>&1
  tail -2 $OUT/plain_$sc.log
done
