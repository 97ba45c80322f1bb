#!/usr/bin/env python3
"""Where an end-to-end frame (bench.py's e2e leg at N = 1) spends its time: parse + compile, upload, rt_render, free."""
import ctypes as C, importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
rt = importlib.import_module("raytracing-project_b200")
path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "scenes", "snorlax_4k.json")
rt.init(0)
sc = rt.load_scene_from_json(path)
host = torch.empty((sc.height, sc.width, 3), dtype=torch.uint8).pin_memory()
acc = {"parse+compile": 0.0, "upload": 0.0, "rt_render": 0.0, "kernel_ms": 0.0, "free": 0.0}
n = 30
for it in range(n + 3):
    t0 = time.perf_counter(); s2 = rt.load_scene_from_json(path)
    t1 = time.perf_counter(); h = s2.device_handle()
    t2 = time.perf_counter(); st = rt.RtStats(); rt._check(rt.lib().rt_render(h, 0, 0, 1, C.c_void_p(host.data_ptr()), None, None, C.byref(st)))
    t3 = time.perf_counter(); s2.close()
    t4 = time.perf_counter()
    if it >= 3:
        for k, v in zip(acc, (t1 - t0, t2 - t1, t3 - t2, st.kernel_ms * 1e-3, t4 - t3)): acc[k] += v
print({k: round(v / n * 1e3, 4) for k, v in acc.items()}, "ms per frame")
