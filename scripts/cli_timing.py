#!/usr/bin/env python3
"""Wall time of the drop-in command line `ray scene.json out.png [--paper]` (fresh process each time: CUDA context,
jitter checkpoint table, render, D2H, PNG)."""
import os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ray = os.path.join(ROOT, "raytracing-project_b200", "ray")
for scene, extra in (("snorlax_4k", []), ("snorlax_4k", []), ("snorlax_4k", []), ("penguin_4k", ["--paper"]), ("pokeballs", [])):
    t = time.perf_counter()
    r = subprocess.run([ray, os.path.join(ROOT, "scenes", scene + ".json"), "/tmp/cli_out.png"] + extra, capture_output=True, text=True)
    dt = time.perf_counter() - t
    print(f"{scene} {' '.join(extra)} rc={r.returncode} wall {dt * 1e3:.0f} ms  png {os.path.getsize('/tmp/cli_out.png')} bytes", flush=True)
