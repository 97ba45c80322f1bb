#!/usr/bin/env python3
"""Wall time of building the jitter checkpoint table (first use in a process): jump-ahead (default) against the
twist-by-twist walk (rt_debug_set_option mt_sequential).  Usage: python scripts/mt_table_timing.py   (spawns one process per case)"""
import importlib, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1:
    sys.path.insert(0, ROOT)
    rt = importlib.import_module("raytracing-project_b200")
    rt.debug_option("mt_sequential", int(sys.argv[2]))
    rt.init(0)
    rt.debug_jitter(0, 8)                                   # context, module load, first small table
    rt.shutdown(); rt.init(0)
    first = int(sys.argv[1])
    t = time.perf_counter(); rt.debug_jitter(first, 8); dt = (time.perf_counter() - t) * 1e3
    print(f"{sys.argv[2]} samples {first + 8} tiles {(first + 8 + 1247) // 1248} ms {dt:.2f}")
else:
    for first in (3840 * 2160 * 8 - 8, 7680 * 4320 * 8 - 8, 16384 * 16384 * 8 - 8):
        for seq in ("1", "0"):
            print(subprocess.run([sys.executable, __file__, str(first), seq], capture_output=True, text=True).stdout.strip(), flush=True)
