#!/usr/bin/env python3
"""The command compute-sanitizer wraps (scripts/gpu_sanitize.sh): small scenes through every kernel variant — per-thread
generic, flat with and without CTA lockstep, wavefront, paper, the global-memory scene path and the known-answer kernel."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rt = importlib.import_module("raytracing-project_b200")
rt.init(0)
scenes = sys.argv[1:] or ["quirks", "synthetic_b_small", "snorlax_small"]
for name in scenes:
    path = os.path.join(ROOT, "scenes", name + ".json")
    for opts in ({}, {"wavefront": 1}, {"wavefront": 0, "flat": 0}, {"cull": 0}, {"flat": 0, "cull": 0}, {"global_scene": 1}):
        for k in rt.DEBUG_OPTIONS:
            rt.debug_option(k, -1)
        for k, v in opts.items():
            rt.debug_option(k, v)
        sc = rt.load_scene_from_json(path)
        for mode in (rt.RenderMode.Standard, rt.RenderMode.Paper):
            tr = rt.Tracer(sc, mode)
            out = tr.render(want_rgb8=True, want_ids=True)
            print(name, opts, mode.name, tr.last_stats["kernel_ms"], int(out["rgb8"].sum()), flush=True)
        rays = np.random.default_rng(1).normal(size=(300, 6))
        rt.Tracer(sc).trace_rays(rays)
        sc.close()
rt.shutdown()
