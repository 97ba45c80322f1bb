#!/bin/bash
# N-GPU session: peer-frame parity check, then bench lines with both exchanges. Usage: bash scripts/gpu_peer.sh <tag> <N> [scenes]
TAG=$1; N=$2; shift 2
OUT=gpurun_out/$TAG; mkdir -p $OUT
SCENES=${@:-"snorlax_4k"}
nvidia-smi -L > $OUT/gpus.txt 2>&1
nvidia-smi topo -m > $OUT/topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $OUT/pytest_multi.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_multi.log
tail -5 $OUT/pytest_multi.log
for sc in $SCENES; do
  for ex in peer gather; do
    f=$OUT/bench_${sc}_n${N}_$ex
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --exchange $ex --scene scenes/$sc.json > $f.json 2> $f.err
    grep '^{' $f.json | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('$sc', '$ex', 'n', j['n_gpus'], 'ms', round(j['ms_per_step'],3), 'kernel_ms', round(j['kernel_ms'],3), 'e2e_ms', round(j['e2e']['ms_per_frame'],2), j['config']['exchange'])" || tail -5 $f.err
  done
done
