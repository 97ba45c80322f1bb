#!/bin/bash
# Round-2 session B: a changed kernel -> smoke (short timeout: a hang must not cost the box), parity suite, bench lines,
# launch list + one full ncu capture of the trace kernel.  Usage under gpurun: bash scripts/gpu_r2b.sh <tag> [scenes for ncu]
TAG=${1:-r2b}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout -k 5 180 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; rc=$?
echo "smoke rc=$rc"; tail -3 $OUT/smoke.log
if [ $rc -ne 0 ]; then echo "smoke failed: stopping"; exit 1; fi
timeout -k 10 1500 python -m pytest tests -m gpu -q -rA --timeout=400 --durations=8 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed" $OUT/pytest_gpu.log | tail -25
timeout -k 10 300 python bench.py --steps 10 --warmup 3 --no-cpu > $OUT/bench_snorlax_4k.json 2> $OUT/bench_snorlax_4k.err
for sc in pokeballs snorlax_4k_reflect synthetic_b_1k synthetic_b_4k synthetic_a_2k synthetic_a_8k penguin_4k; do
  timeout -k 10 300 python bench.py --steps 3 --warmup 3 --no-cpu --scene scenes/$sc.json > $OUT/bench_$sc.json 2> $OUT/bench_$sc.err
done
for sc in penguin penguin_4k snorlax_4k; do
  timeout -k 10 300 python bench.py --steps 5 --warmup 3 --no-cpu --paper --scene scenes/$sc.json > $OUT/bench_${sc}_paper.json 2> $OUT/bench_${sc}_paper.err
done
python - <<PY
import json, glob, os
for f in sorted(glob.glob("$OUT/bench_*.json")):
    try:
        d = json.load(open(f))
        print(os.path.basename(f), "ms", round(d["ms_per_frame"], 3), "e2e", round(d["e2e"]["ms_per_frame"], 3), "tests/ray", round(d["rays"]["prim_tests"] / d["rays_per_frame"], 2))
    except Exception as e:
        print(os.path.basename(f), "ERR", e)
PY
for sc in ${2:-snorlax_4k}; do
timeout -k 10 300 python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/plain_prof_$sc.log 2>&1 &&
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k "regex:render_(standard|wavefront)" -s 1 -c 1 -o $OUT/prof_$sc \
    python scripts/profile_run.py scenes/$sc.json standard 2 > $OUT/ncu_prof_$sc.log 2>&1
done
