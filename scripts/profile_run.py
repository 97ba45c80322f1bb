#!/usr/bin/env python3
"""Render one scene a few times through the C ABI (the command ncu wraps)."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("raytracing-project_b200")
path = sys.argv[1]; mode = rt.RenderMode.Paper if len(sys.argv) > 2 and sys.argv[2] == "paper" else rt.RenderMode.Standard
n = int(sys.argv[3]) if len(sys.argv) > 3 else 2
rt.init(0)
for kv in os.environ.get("RT_OPT", "").split(","):      # e.g. RT_OPT=cull=0,flat=0 (rt_debug_set_option)
    if "=" in kv:
        rt.debug_option(kv.split("=")[0], int(kv.split("=")[1]))
tr = rt.Tracer(rt.load_scene_from_json(path), mode)
import zlib
for _ in range(n):
    out = tr.render(want_rgb8=True)
    st = dict(tr.last_stats); st["rgb8_crc32"] = zlib.crc32(out["rgb8"].tobytes())      # frame identity across variants
    print(st)
