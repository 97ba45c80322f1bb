#!/usr/bin/env python3
"""Render one scene a few times through the C ABI (the command ncu wraps)."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("raytracing-project_b200")
path = sys.argv[1]; mode = rt.RenderMode.Paper if len(sys.argv) > 2 and sys.argv[2] == "paper" else rt.RenderMode.Standard
n = int(sys.argv[3]) if len(sys.argv) > 3 else 2
rt.init(0)
tr = rt.Tracer(rt.load_scene_from_json(path), mode)
for _ in range(n):
    tr.render(want_rgb8=True)
    print(tr.last_stats)
