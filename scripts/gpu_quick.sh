#!/bin/bash
# Quick GPU iteration: parity suite + kernel timings of the bench scenes (no CPU leg, no ncu).
# Usage: bash scripts/gpu_quick.sh <tag> [scenes...]
TAG=${1:-q}; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
SCENES=${@:-"snorlax_4k snorlax_4k_reflect pokeballs synthetic_b_4k synthetic_a_8k"}
for sc in $SCENES; do
  python bench.py --steps 5 --warmup 3 --no-cpu --scene scenes/$sc.json > $OUT/bench_$sc.json 2> $OUT/bench_$sc.err
done
python bench.py --steps 5 --warmup 3 --no-cpu --paper --scene scenes/penguin_4k.json > $OUT/bench_penguin4k_paper.json 2> $OUT/bench_penguin4k_paper.err
tail -4 $OUT/pytest_gpu.log
python - <<PY
import json,glob
for f in sorted(glob.glob("$OUT/bench_*.json")):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], "ms", round(j["ms_per_step"],3), "Mrays/s", round(j["value"]), "frac", round(j["roofline"]["frac"],4), "e2e_ms", round(j["e2e"]["ms_per_frame"],2), "tests", j["rays"]["prim_tests"])
    except Exception as e:
        print(f, "ERR", e)
PY
