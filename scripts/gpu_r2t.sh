#!/bin/bash
# Grid on / off: kernel ms (+ the grid identity test).
TAG=${1:-r2t}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s --timeout=400 -k "grid_does_not or wavefront_and or scene_matches_oracle" > $OUT/pytest_subset.log 2>&1; echo "subset rc=$?"
grep -E "with the grid|passed|failed|Error" $OUT/pytest_subset.log | tail -8
for opt in ${3:-"" "grid=0"}; do
  for a in ${2:-synthetic_b_1k synthetic_b_4k}; do
    RT_OPT=$opt timeout -k 10 300 python scripts/profile_run.py scenes/$a.json standard 3 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read().strip()); print('[$opt]','$a','kernel_ms',round(d['kernel_ms'],3),'tests/ray',round(d['prim_tests']/(d['primary_rays']+d['shadow_rays']+d['secondary_rays']),2))" | tee -a $OUT/ab.txt
  done
done
