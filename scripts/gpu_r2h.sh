#!/bin/bash
# Session r2h: host-direct output / per-device contexts: new tests first, then the whole suite, then the bench line.
TAG=${1:-r2h}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout -k 10 600 python -m pytest tests/test_gpu_host_frame.py tests/test_zz_no_alloc.py -m gpu -x -q -rA --timeout=300 > $OUT/pytest_host_frame.log 2>&1; echo "host_frame rc=$?"
grep -E "^FAILED|^ERROR|passed|failed|Error|error" $OUT/pytest_host_frame.log | tail -15
timeout -k 10 1500 python -m pytest tests -m gpu -q -rA --timeout=400 --durations=8 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed" $OUT/pytest_gpu.log | tail -15
timeout -k 10 600 python bench.py --steps 10 --warmup 3 > $OUT/bench_snorlax_4k.json 2> $OUT/bench_snorlax_4k.err; cat $OUT/bench_snorlax_4k.json; tail -3 $OUT/bench_snorlax_4k.err
timeout -k 10 600 python bench.py --steps 5 --warmup 3 --paper --scene scenes/penguin_4k.json > $OUT/bench_penguin_4k_paper.json 2> $OUT/bench_penguin_4k_paper.err; cat $OUT/bench_penguin_4k_paper.json; tail -3 $OUT/bench_penguin_4k_paper.err
python scripts/cli_timing.py > $OUT/cli_timing.txt 2>&1; tail -5 $OUT/cli_timing.txt
