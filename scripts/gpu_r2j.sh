#!/bin/bash
# A/B of library variants + the parity subset that exercises the masked kernels.  Usage: gpu_r2j.sh <tag> "<variants>" ["<scenes>"]
TAG=${1:-r2j}; OUT=gpurun_out/$TAG; mkdir -p $OUT
SC=${3:-"snorlax_4k snorlax_4k_reflect penguin_4k pokeballs synthetic_a_2k synthetic_b_1k penguin_4k:paper snorlax_4k:paper"}
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_host_frame.py -m gpu -x -q --timeout=400 -k "not bands and not native and not jitter and not checkpoint" > $OUT/pytest_subset.log 2>&1; echo "parity subset rc=$?"
tail -3 $OUT/pytest_subset.log
bash scripts/gpu_ab.sh $TAG "$SC" "$2"
timeout -k 10 300 python scripts/e2e_breakdown.py 2>&1 | tail -1 | tee $OUT/e2e_breakdown.txt
