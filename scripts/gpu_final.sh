#!/bin/bash
# Final single-GPU session of a round: smoke, the whole GPU test suite, the headline bench line, one bench line per
# BASELINE config with its cpu_baseline, the ncu launch list of the bench command and one `ncu --set full` capture per
# kernel family (each after a plain run of the same command exited 0).   Usage under gpurun: bash scripts/gpu_final.sh <tag>
TAG=${1:-final}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log
timeout -k 10 1200 python -m pytest tests -m gpu -x -q -rA --timeout=600 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
timeout -k 10 600 python bench.py --steps 10 --warmup 3 > $OUT/bench_snorlax_4k.json 2> $OUT/bench_snorlax_4k.err
for sc in pokeballs snorlax_4k_reflect synthetic_a_8k synthetic_b_4k; do
  timeout -k 10 600 python bench.py --steps 3 --warmup 3 --cpu-seconds 6 --scene scenes/$sc.json > $OUT/bench_$sc.json 2> $OUT/bench_$sc.err
done
timeout -k 10 600 python bench.py --steps 5 --warmup 3 --paper --scene scenes/penguin.json > $OUT/bench_penguin_paper.json 2> $OUT/bench_penguin_paper.err
timeout -k 10 600 python bench.py --steps 5 --warmup 3 --paper --scene scenes/penguin_4k.json > $OUT/bench_penguin_4k_paper.json 2> $OUT/bench_penguin_4k_paper.err
for f in $OUT/bench_*.json; do grep '^{' $f | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); c=j.get('cpu_baseline',{}); print(j['config']['workload'], j['config']['mode'], 'ms', round(j['ms_per_step'],3), 'e2e', round(j['e2e']['ms_per_frame'],3), 'Mrays/s', round(j['value']), 'frac', round(j['roofline']['frac'],3), 'cpu', c.get('value'))" || echo "FAILED $f"; done
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
ls $OUT | tr '\n' ' '
timeout -k 10 600 python bench.py --steps 2 --warmup 3 --no-cpu --scene scenes/synthetic_b_16k.json > $OUT/bench_synthetic_b_16k.json 2> $OUT/bench_synthetic_b_16k.err; grep -o '"ms_per_step": [0-9.]*' $OUT/bench_synthetic_b_16k.json
