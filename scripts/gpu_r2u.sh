#!/bin/bash
TAG=${1:-r2u}; OUT=gpurun_out/$TAG; mkdir -p $OUT
sc=synthetic_b_1k
timeout -k 10 300 python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/plain_prof_$sc.log 2>&1 &&
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k "regex:render_(standard|wavefront)" -s 1 -c 1 -o $OUT/prof_$sc \
    python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/ncu_prof_$sc.log 2>&1
tail -1 $OUT/plain_prof_$sc.log
