#!/bin/bash
# One GPU-box session: parity tests, bench lines, ncu launch list + full capture of the trace kernel.
# Usage (from the repo root, under gpurun):  bash scripts/gpu_round.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > $OUT/bench_snorlax4k.json 2> $OUT/bench_snorlax4k.err
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err
for sc in pokeballs snorlax_4k_reflect synthetic_b_4k synthetic_a_8k; do
  python bench.py --steps 3 --warmup 3 --no-cpu --scene scenes/$sc.json > $OUT/bench_$sc.json 2> $OUT/bench_$sc.err
done
python bench.py --steps 5 --warmup 3 --no-cpu --paper --scene scenes/penguin_4k.json > $OUT/bench_penguin4k_paper.json 2> $OUT/bench_penguin4k_paper.err
python scripts/gpu_parity.py > $OUT/parity.log 2>&1
# ncu: launch list of the bench command, then one full capture of the trace kernel (after a plain run exited 0)
python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/ncu_bench.log 2>&1
for sc in snorlax_4k synthetic_a_2k synthetic_b_1k; do
python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/plain_prof_$sc.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:render_(standard|wavefront)" -s 1 -c 1 -o $OUT/prof_$sc \
    python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/ncu_prof_$sc.log 2>&1
done
python scripts/profile_run.py scenes/penguin_4k.json paper 2 > $OUT/plain_prof_paper.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:paper_ -s 2 -c 2 -o $OUT/prof_penguin_4k_paper \
    python scripts/profile_run.py scenes/penguin_4k.json paper 2 > $OUT/ncu_prof_paper.log 2>&1
tail -3 $OUT/pytest_gpu.log; cat $OUT/bench_snorlax4k.json
