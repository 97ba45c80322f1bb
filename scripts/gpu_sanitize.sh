#!/bin/bash
# compute-sanitizer over the small scenes (SURVEY.md §5 race-detection row).  Usage under gpurun: bash scripts/gpu_sanitize.sh <tag>
TAG=${1:-san}; OUT=gpurun_out/$TAG; mkdir -p $OUT
python scripts/sanitize_run.py quirks > $OUT/plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/plain.log; exit 1; }
for tool in memcheck racecheck synccheck initcheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_run.py quirks synthetic_b_small snorlax_small > $OUT/$tool.log 2>&1
  echo "$tool rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $OUT/$tool.log | tail -1)"
done
