#!/bin/bash
# Profile session on the current build: bench line, ncu launch list, ncu --set full of each kernel family.
# Usage under gpurun: bash scripts/gpu_r2m.sh [tag]
TAG=${1:-r2m}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout -k 10 600 python bench.py --steps 10 --warmup 3 > $OUT/bench_snorlax_4k.json 2> $OUT/bench_snorlax_4k.err
timeout -k 10 300 python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/plain_bench.log 2>&1 &&
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/ncu_bench.log 2>&1
for sc in snorlax_4k synthetic_a_2k synthetic_b_1k; do
timeout -k 10 300 python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/plain_prof_$sc.log 2>&1 &&
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k "regex:render_(standard|wavefront)" -s 1 -c 1 -o $OUT/prof_$sc \
    python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/ncu_prof_$sc.log 2>&1
done
timeout -k 10 300 python scripts/profile_run.py scenes/penguin_4k.json paper 2 > $OUT/plain_prof_paper.log 2>&1 &&
timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:paper_ -s 2 -c 2 -o $OUT/prof_penguin_4k_paper \
    python scripts/profile_run.py scenes/penguin_4k.json paper 2 > $OUT/ncu_prof_paper.log 2>&1
cat $OUT/bench_snorlax_4k.json; ls -la $OUT
