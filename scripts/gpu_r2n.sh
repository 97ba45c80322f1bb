#!/bin/bash
# A/B of library variants + the parity subset (standard, paper, host frame, no-alloc).  Usage: gpu_r2n.sh <tag> "<variants>" ["<scenes>"]
TAG=${1:-r2n}; OUT=gpurun_out/$TAG; mkdir -p $OUT
SC=${3:-"snorlax_4k snorlax_4k_reflect penguin_4k synthetic_a_2k synthetic_b_1k penguin_4k:paper snorlax_4k:paper"}
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_host_frame.py tests/test_zz_no_alloc.py -m gpu -x -q --timeout=400 -k "not bands and not native and not jitter and not checkpoint" > $OUT/pytest_subset.log 2>&1; echo "parity subset rc=$?"
tail -3 $OUT/pytest_subset.log
bash scripts/gpu_ab.sh $TAG "$SC" "$2"
