#!/bin/bash
# Parity (incl. the jitter-stream tests) + kernel ms + frame checksums.
TAG=${1:-r2w}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_host_frame.py tests/test_zz_no_alloc.py tests/test_reference_kats.py -m gpu -x -q --timeout=400 -k "not bands and not native" > $OUT/pytest_subset.log 2>&1; echo "parity subset rc=$?"; tail -2 $OUT/pytest_subset.log
for a in snorlax_4k snorlax_4k_reflect penguin_4k pokeballs synthetic_a_2k synthetic_b_1k penguin_4k:paper; do
  sc=${a%%:*}; mode=standard; [[ "$a" == *:* ]] && mode=${a##*:}
  timeout -k 10 300 python scripts/profile_run.py scenes/$sc.json $mode 4 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read().strip()); print('$a','kernel_ms',round(d['kernel_ms'],3),'tests/ray',round(d['prim_tests']/(d['primary_rays']+d['shadow_rays']+d['secondary_rays']),2),'crc',d['rgb8_crc32'])" | tee -a $OUT/ab.txt
done
RT_OPT=grid=0 timeout -k 10 300 python scripts/profile_run.py scenes/synthetic_b_1k.json standard 2 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read().strip()); print('[grid=0] synthetic_b_1k','kernel_ms',round(d['kernel_ms'],3),'rays',d['shadow_rays'],d['secondary_rays'],'crc',d['rgb8_crc32'])" | tee -a $OUT/ab.txt
timeout -k 10 300 python scripts/profile_run.py scenes/synthetic_b_1k.json standard 2 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read().strip()); print('[grid] synthetic_b_1k','kernel_ms',round(d['kernel_ms'],3),'rays',d['shadow_rays'],d['secondary_rays'],'crc',d['rgb8_crc32'])" | tee -a $OUT/ab.txt
