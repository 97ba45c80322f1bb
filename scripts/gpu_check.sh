#!/bin/bash
# Last check of a build: smoke, the whole GPU suite, the default bench line.
TAG=${1:-chk}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log
timeout -k 10 1200 python -m pytest tests -m gpu -x -q -rA --timeout=600 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -2 $OUT/pytest_gpu.log
timeout -k 10 600 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; tail -c 1500 $OUT/bench_default.json | head -c 700
