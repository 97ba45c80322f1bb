#!/bin/bash
# Multi-GPU bench lines (one process per GPU, torchrun).
# Usage: bash scripts/gpu_multi.sh <tag> <scene>:<n1>,<n2>,...[:steps] ...     e.g.  m8 snorlax_4k:1,2,4,8 synthetic_b_16k:8:2
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt 2>&1
for tok in "$@"; do
  IFS=: read sc ns steps <<< "$tok"; steps=${steps:-5}
  for n in ${ns//,/ }; do
    f=$OUT/bench_${sc}_n$n
    if [ $n -eq 1 ]; then
      timeout 600 python bench.py --gpus 1 --steps $steps --warmup 3 --no-cpu --scene scenes/$sc.json > $f.json 2> $f.err
    else
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $n --steps $steps --warmup 3 --no-cpu --scene scenes/$sc.json > $f.json 2> $f.err
    fi
    grep '^{' $f.json | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('$sc', 'n', j['n_gpus'], 'ms', round(j['ms_per_step'],3), 'kernel_ms', round(j['kernel_ms'],3), 'e2e_ms', round(j['e2e']['ms_per_frame'],2), 'Mrays/s', round(j['value']))" || tail -5 $f.err
  done
done
