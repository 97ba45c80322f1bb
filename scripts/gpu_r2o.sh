#!/bin/bash
# Parity subset + A/B of library variants + one ncu --set full capture of the snorlax trace kernel (default build).
TAG=${1:-r2o}; OUT=gpurun_out/$TAG; mkdir -p $OUT
bash scripts/gpu_r2n.sh $TAG "$2" "$3"
timeout -k 10 300 python scripts/profile_run.py scenes/snorlax_4k.json standard 2 > $OUT/plain_prof_snorlax_4k.log 2>&1 &&
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k "regex:render_(standard|wavefront)" -s 1 -c 1 -o $OUT/prof_snorlax_4k \
    python scripts/profile_run.py scenes/snorlax_4k.json standard 2 > $OUT/ncu_prof_snorlax_4k.log 2>&1
ls $OUT
