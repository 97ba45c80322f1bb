#!/usr/bin/env python3
"""Quick GPU-vs-oracle parity report over scene files (development aid; the gate is tests/ -m gpu)."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
rt = importlib.import_module("raytracing-project_b200")
import oracle_api as oa

def compare(path, mode):
    sc = rt.load_scene_from_json(path)
    tr = rt.Tracer(sc, rt.RenderMode.Paper if mode == "paper" else rt.RenderMode.Standard)
    g = tr.render(want_rgb8=True, want_ids=True)
    t0 = time.time(); g = tr.render(want_rgb8=True, want_ids=True); tg = time.time() - t0
    st = tr.last_stats
    t0 = time.time(); o = oa.Oracle.from_file(path).render(mode, want_ids=True); to = time.time() - t0
    ob = oa.to_byte(o["fb"])
    diff = np.abs(g["rgb8"].astype(int) - ob.astype(int)).max(axis=2)
    id_bad = (g["ids"] != o["ids"])
    rec = {"scene": os.path.basename(path), "mode": mode, "size": [sc.width, sc.height],
           "kernel_ms": round(st["kernel_ms"], 3), "gpu_call_s": round(tg, 4), "oracle_s": round(to, 2),
           "px_within_1": float((diff <= 1).mean()), "px_exact": float((diff == 0).mean()), "max_diff": int(diff.max()),
           "id_mismatch_samples": int(id_bad.sum()), "id_mismatch_frac": float(id_bad.mean()),
           "rays": {k: st[k] for k in ("primary_rays", "shadow_rays", "secondary_rays", "prim_tests")}}
    print(json.dumps(rec), flush=True)
    return rec

if __name__ == "__main__":
    rt.init(0)
    print("fp32 peak TFLOP/s", rt.measure_fp32_peak(), flush=True)
    j = rt.debug_jitter(0, 4000); oj = oa.jitter(0, 4000)
    print("jitter first", j[0], "equal:", np.array_equal(j, oj), flush=True)
    j = rt.debug_jitter(123456789, 3000); oj = oa.jitter(123456789, 3000)
    print("jitter far equal:", np.array_equal(j, oj), flush=True)
    args = sys.argv[1:] or ["quirks:standard", "quirks:paper", "pokeballs_small:standard", "penguin_small:standard",
                            "penguin_small:paper", "snorlax_small:standard", "snorlax_4k_reflect_small:standard",
                            "synthetic_a_tiny:standard", "synthetic_b_small:standard"]
    for a in args:
        name, mode = a.split(":")
        compare(os.path.join(ROOT, "scenes", name + ".json"), mode)
