"""bench.py prints the same `config` object from both arms (ours and --impl reference); the reference arm derives it from the
scene file alone.  It must agree with what the scene compiler reports for every scene in scenes/."""
import glob
import importlib.util
import os

from conftest import ROOT


def test_workload_config_matches_the_compiled_scene(rt):
    spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec); spec.loader.exec_module(bench)
    paths = sorted(p for p in glob.glob(os.path.join(ROOT, "scenes", "*.json")) if "16k" not in p and "8k" not in p)
    assert len(paths) >= 15
    for p in paths:
        c = bench.workload_config(p, False, 1)
        s = rt.load_scene_from_json(p); d = s.desc
        assert (c["width"], c["height"], c["recursion"], c["lights"], c["objects"], c["csg_leaves"]) == \
               (s.width, s.height, int(d.recursion), int(d.n_lights), int(d.n_objects), int(d.n_leaves)), p
        assert bench.workload_config(p, False, 8)["parallelism"] == "row-tile interleave x8"
        assert set(c) == set(bench.workload_config(p, True, 2))          # same keys whatever the mode / world
