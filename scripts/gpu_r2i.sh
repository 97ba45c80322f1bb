#!/bin/bash
# Session r2i (N GPUs of one box): multi-device tests, torchrun bench at N, RAY_GPUS CLI.  Usage: gpu_r2i.sh <tag> <N>
TAG=${1:-r2i}; N=${2:-2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi -L > $OUT/smi.txt 2>&1; df -h /dev/shm >> $OUT/smi.txt 2>&1
timeout -k 10 900 python -m pytest tests/test_gpu_host_frame.py tests/test_gpu_multi.py -m gpu -q -rA --timeout=600 > $OUT/pytest_multi.log 2>&1; echo "multi tests rc=$?"
grep -E "^FAILED|^ERROR|passed|failed|SKIPPED" $OUT/pytest_multi.log | tail -15
for n in 1 $N; do
  if [ $n = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29655"; fi
  timeout -k 10 600 $L bench.py --gpus $n --steps 10 --warmup 3 --no-cpu > $OUT/bench_snorlax_4k_n$n.json 2> $OUT/bench_snorlax_4k_n$n.err
  python - <<PY
import json
try:
    d = json.loads(open("$OUT/bench_snorlax_4k_n$n.json").read().strip().splitlines()[-1])
    print("N=$n ms_per_frame %.3f kernel %.3f e2e %.3f path=%s equal=%s" % (d["ms_per_frame"], d["kernel_ms"], d["e2e"]["ms_per_frame"], d["e2e"].get("path"), d.get("frame_equals_single_gpu", d.get("host_frame_equals_device_frame"))))
except Exception as e:
    print("N=$n bench failed", e); print(open("$OUT/bench_snorlax_4k_n$n.err").read()[-1500:])
PY
done
for g in 1 $N; do
  for i in 1 2; do
  t0=$(date +%s.%N); RAY_GPUS=$g RAY_STATS=1 raytracing-project_b200/ray scenes/snorlax_4k.json /tmp/o_$g.png 2>&1 | tail -1; t1=$(date +%s.%N)
  echo "RAY_GPUS=$g wall $(echo "$t1 - $t0" | bc) s"
  done
done 2>&1 | tee $OUT/cli_gpus.txt
cmp /tmp/o_1.png /tmp/o_$N.png && echo "PNG identical 1 vs $N GPUs" | tee -a $OUT/cli_gpus.txt
