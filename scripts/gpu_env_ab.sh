#!/bin/bash
# A/B of a run-time switch (environment variable read at scene upload) on kernel ms.
# Usage: bash scripts/gpu_env_ab.sh <tag> <VAR> "<values>" "<scenes>"
TAG=$1; VAR=$2; VALS=$3; SCENES=$4
OUT=gpurun_out/$TAG; mkdir -p $OUT
for sc in $SCENES; do
  for v in $VALS; do
    env $VAR=$v python scripts/profile_run.py scenes/$sc.json standard 4 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read().strip()); print('$sc','$VAR=$v','kernel_ms',round(d['kernel_ms'],3))" | tee -a $OUT/ab.txt
  done
done
