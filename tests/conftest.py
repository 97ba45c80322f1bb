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
