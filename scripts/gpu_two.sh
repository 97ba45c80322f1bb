#!/bin/bash
# Two-GPU session: the tests that need two devices, then bench lines at N = 2.
TAG=${1:-g2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout -k 10 600 python -m pytest tests/test_gpu_host_frame.py tests/test_gpu_multi.py -m gpu -q -rA --timeout=400 > $OUT/two_gpu_tests.log 2>&1; echo "two-gpu tests rc=$?"; tail -2 $OUT/two_gpu_tests.log
bash scripts/gpu_multi.sh $TAG snorlax_4k:2:10 synthetic_b_16k:2:2
