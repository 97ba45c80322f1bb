#!/bin/bash
# Round-2 session A: the widened parity net on the kernels as they are + a bench line with cpu_baseline for every config.
TAG=${1:-r2a}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
nproc > $OUT/nproc.txt
timeout 1500 python -m pytest tests -m gpu -q -rA --durations=15 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
grep -E "passed|failed" $OUT/pytest_gpu.log | tail -2
python bench.py --steps 10 --warmup 3 --cpu-seconds 8 > $OUT/bench_snorlax_4k.json 2> $OUT/bench_snorlax_4k.err
for sc in pokeballs snorlax_4k_reflect synthetic_b_4k synthetic_a_8k; do
  timeout 600 python bench.py --steps 3 --warmup 3 --cpu-seconds 8 --scene scenes/$sc.json > $OUT/bench_$sc.json 2> $OUT/bench_$sc.err
done
for sc in penguin penguin_4k; do
  timeout 600 python bench.py --steps 5 --warmup 3 --paper --scene scenes/$sc.json > $OUT/bench_${sc}_paper.json 2> $OUT/bench_${sc}_paper.err
done
bash scripts/gpu_sanitize.sh $TAG/sanitizer
cat $OUT/bench_snorlax_4k.json | cut -c1-600
