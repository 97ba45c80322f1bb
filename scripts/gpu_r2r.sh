#!/bin/bash
# A/B of variants on the parity subset's scenes, the parity subset under the variant too, then configs[4] at N = 1.
TAG=${1:-r2r}; OUT=gpurun_out/$TAG; mkdir -p $OUT
bash scripts/gpu_ab.sh $TAG "snorlax_4k snorlax_4k_reflect penguin_4k pokeballs synthetic_b_1k penguin_4k:paper" "$2"
for v in $3; do
  cp raytracing-project_b200/librt_b200.so /tmp/default.so; cp raytracing-project_b200/variants/$v.so raytracing-project_b200/librt_b200.so
  timeout -k 10 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_host_frame.py -m gpu -x -q --timeout=400 -k "not bands and not native and not jitter and not checkpoint" > $OUT/pytest_subset_$v.log 2>&1; echo "parity subset under $v rc=$?"; tail -1 $OUT/pytest_subset_$v.log
  cp /tmp/default.so raytracing-project_b200/librt_b200.so
done
timeout -k 10 600 python bench.py --steps 2 --warmup 3 --no-cpu --scene scenes/synthetic_b_16k.json > $OUT/bench_synthetic_b_16k_n1.json 2> $OUT/bench_synthetic_b_16k_n1.err; tail -c 600 $OUT/bench_synthetic_b_16k_n1.json | head -c 300
