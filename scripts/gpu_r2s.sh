#!/bin/bash
# Grid for incoherent rays: parity subset (incl. the grid identity test), kernel ms with the grid on / off.
TAG=${1:-r2s}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_host_frame.py tests/test_zz_no_alloc.py tests/test_reference_kats.py -m gpu -x -q -s --timeout=400 -k "not bands and not native and not jitter and not checkpoint" > $OUT/pytest_subset.log 2>&1; echo "parity subset rc=$?"
grep -E "with the grid|passed|failed|Error" $OUT/pytest_subset.log | tail -12
for opt in "" "grid=0"; do
  for a in synthetic_b_1k synthetic_b_4k snorlax_4k_reflect snorlax_4k synthetic_a_2k; do
    RT_OPT=$opt timeout -k 10 300 python scripts/profile_run.py scenes/$a.json standard 3 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read().strip()); print('[$opt]','$a','kernel_ms',round(d['kernel_ms'],3),'tests/ray',round(d['prim_tests']/(d['primary_rays']+d['shadow_rays']+d['secondary_rays']),2))" | tee -a $OUT/ab.txt
  done
done
