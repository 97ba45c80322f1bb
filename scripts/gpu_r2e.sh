#!/bin/bash
# smoke + parity suite + A/B timings + optional ncu captures.  Usage: gpu_r2e.sh <tag> "<variants>" "<ncu scenes>"
TAG=${1:-r2e}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout -k 5 180 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; rc=$?
echo "smoke rc=$rc"; tail -2 $OUT/smoke.log
if [ $rc -ne 0 ]; then echo "smoke failed: stopping"; exit 1; fi
timeout -k 10 1500 python -m pytest tests -m gpu -q -rA --timeout=400 --durations=5 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed" $OUT/pytest_gpu.log | tail -25
bash scripts/gpu_ab.sh $TAG "snorlax_4k snorlax_4k_reflect penguin_4k pokeballs synthetic_a_2k synthetic_b_1k penguin_4k:paper snorlax_4k:paper" "$2"
for sc in $3; do
timeout -k 10 300 python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/plain_prof_$sc.log 2>&1 &&
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k "regex:render_(standard|wavefront)" -s 1 -c 1 -o $OUT/prof_$sc \
    python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/ncu_prof_$sc.log 2>&1
done
