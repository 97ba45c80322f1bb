import glob
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(ROOT, "scenes")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def rt():
    """The product package (ctypes over librt_b200.so).  Built in-tree; never falls back to anything."""
    mod = importlib.import_module("raytracing-project_b200")
    if not os.path.exists(mod.LIB_PATH):
        mod.build()
    return mod


@pytest.fixture(scope="session")
def oa():
    import oracle_api
    oracle_api.lib()
    return oracle_api


@pytest.fixture(scope="session")
def gpu(rt):
    rt.init(0)
    return rt


@pytest.fixture
def force(gpu):
    """force(name, value): rt_debug_set_option for this test only (all hooks return to the library default afterwards)."""
    yield gpu.debug_option
    for name in gpu.DEBUG_OPTIONS:
        gpu.debug_option(name, -1)


def golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


@pytest.fixture(scope="session")
def goldens():
    out = {}
    for n in golden_names():
        z = np.load(os.path.join(GOLDEN, n + ".npz"))
        out[n] = {k: z[k] for k in z.files}
        out[n]["scene_json"] = str(out[n]["scene_json"])
    return out


def scene_path(name: str) -> str:
    return os.path.join(SCENES, name if name.endswith(".json") else name + ".json")


def image_parity(rgb8_a: np.ndarray, rgb8_b: np.ndarray):
    """Fraction of pixels whose channels all agree within 1/255, and the largest channel difference."""
    d = np.abs(rgb8_a.astype(int) - rgb8_b.astype(int)).max(axis=2)
    return float((d <= 1).mean()), int(d.max())


def golden_oracle(oa, g):
    """The oracle's scene for a golden fixture (JSON + the DirectionalLights the fixture was rendered with, if any)."""
    o = oa.Oracle(g["scene_json"])
    for row in g.get("dir_lights", []):
        o.add_dir_light(row[:3], row[3:])
    return o


def golden_gpu_scene(gpu, g):
    sc = gpu.load_scene_from_json_text(g["scene_json"])
    for row in g.get("dir_lights", []):
        sc.add_dir_light(row[:3], row[3:])
    return sc
