#!/bin/bash
# Multi-GPU bench lines (one process per GPU, torchrun). Usage: bash scripts/gpu_multi.sh <tag> <N> [scenes...]
TAG=$1; N=$2; shift 2
OUT=gpurun_out/$TAG; mkdir -p $OUT
SCENES=${@:-"snorlax_4k synthetic_b_4k"}
nvidia-smi -L > $OUT/gpus.txt 2>&1
for sc in $SCENES; do
  for n in 1 $N; do
    if [ $n -eq 1 ]; then
      python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu --scene scenes/$sc.json > $OUT/bench_${sc}_n1.json 2> $OUT/bench_${sc}_n1.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $n --steps 5 --warmup 3 --no-cpu --scene scenes/$sc.json > $OUT/bench_${sc}_n$n.json 2> $OUT/bench_${sc}_n$n.err
    fi
    tail -1 $OUT/bench_${sc}_n$n.json | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('$sc', 'n', j['n_gpus'], 'ms', round(j['ms_per_step'],3), 'kernel_ms', round(j['kernel_ms'],3), 'e2e_ms', round(j['e2e']['ms_per_frame'],2), 'Mrays/s', round(j['value']))" || tail -5 $OUT/bench_${sc}_n$n.err
  done
done
