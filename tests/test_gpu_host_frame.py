"""Host-direct output and the in-process multi-GPU path (include/rt_b200.h: rt_host_frame_*, rt_render_multi).

The render kernels store the packed RGB8 frame straight into pinned, device-mapped host memory, tile by tile; with several
ranks (processes or devices of one process) every rank fills its own tiles of ONE frame.  Whatever the route, the bytes
must equal the frame of the staged path (device frame + pack + copy), which test_gpu_parity.py holds against the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import scene_path

pytestmark = pytest.mark.gpu
MODES = {"standard": 0, "paper": 1}


def _gpus() -> int:
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def staged(gpu, scene, mode):
    return gpu.Tracer(scene, gpu.RenderMode(MODES[mode])).render(want_rgb8=True)["rgb8"]


def direct(gpu, scene, mode, frame, rank=0, world=1):
    import ctypes as C
    gpu._check(gpu.lib().rt_render(scene.device_handle(), MODES[mode], rank, world, C.c_void_p(frame.ptr), None, None, None))


def odd_width_scene(gpu, name, width):
    """The scene at a resolution whose row length is not a multiple of 4 (the word path of the tile flush falls back to bytes)."""
    js = json.load(open(scene_path(name)))
    scr = js["screen"]                                    # nx = round(Lx * dpi), dpi an integer (json_loader.cpp:157, camera.h:19)
    scr["dpi"] = max(1, round(width / scr["dimensions"][0]))
    scr["dimensions"][0] = (width + 0.01) / scr["dpi"]
    sc = gpu.load_scene_from_json_text(json.dumps(js))
    assert sc.width == width, sc.width
    return sc


@pytest.mark.parametrize("mode", ["standard", "paper"])
@pytest.mark.parametrize("name", ["quirks", "pokeballs_small", "synthetic_b_small", "synthetic_a_tiny", "snorlax_small"])
def test_host_direct_frame_equals_staged_frame(gpu, name, mode):
    sc = gpu.load_scene_from_json(scene_path(name))
    want = staged(gpu, sc, mode)
    frame = gpu.HostFrame(sc.height, sc.width)
    frame.array[:] = 7
    direct(gpu, sc, mode, frame)
    assert np.array_equal(frame.array, want), f"{(frame.array != want).any(axis=2).sum()} pixels differ"
    frame.close()


@pytest.mark.parametrize("mode", ["standard", "paper"])
@pytest.mark.parametrize("width", [157, 201, 322])
def test_host_direct_rows_not_a_multiple_of_four(gpu, mode, width):
    sc = odd_width_scene(gpu, "quirks", width)
    want = staged(gpu, sc, mode)
    frame = gpu.HostFrame(sc.height, sc.width)
    direct(gpu, sc, mode, frame)
    assert np.array_equal(frame.array, want)
    frame.close()


@pytest.mark.parametrize("mode", ["standard", "paper"])
@pytest.mark.parametrize("world", [2, 5])
def test_ranks_fill_one_host_frame(gpu, mode, world):
    """Ranks rendered one after the other on one GPU: every rank writes its own tiles only, together the whole frame."""
    sc = odd_width_scene(gpu, "quirks", 250) if world == 5 else gpu.load_scene_from_json(scene_path("quirks"))
    want = staged(gpu, sc, mode)
    frame = gpu.HostFrame(sc.height, sc.width)
    frame.array[:] = 0
    seen = np.zeros(frame.array.shape[:2], np.int32)
    for rank in range(world):
        before = frame.array.copy()
        direct(gpu, sc, mode, frame, rank, world)
        touched = (frame.array != before).any(axis=2)
        assert not (touched & (seen > 0)).any(), "a rank overwrote pixels of another rank"
        seen += touched
    assert np.array_equal(frame.array, want)
    frame.close()


def test_shared_memory_host_frame_two_mappings(gpu):
    """What two processes do, inside one: the creator's mapping and a second rt_host_frame_open mapping of the same POSIX-shm
    name; rank 0 renders through one, rank 1 through the other, and the creator sees the whole frame."""
    sc = gpu.load_scene_from_json(scene_path("pokeballs_small"))
    want = staged(gpu, sc, "standard")
    name = "/rt_b200_test_%d" % os.getpid()
    a = gpu.HostFrame(sc.height, sc.width, name=name, create=True)
    b = gpu.HostFrame(sc.height, sc.width, name=name, create=False)
    with pytest.raises(gpu.RtError):
        gpu.HostFrame(sc.height, sc.width, name=name, create=True)          # the name exists
    a.array[:] = 0
    direct(gpu, sc, "standard", a, 0, 2)
    direct(gpu, sc, "standard", b, 1, 2)
    assert np.array_equal(a.array, want) and np.array_equal(b.array, want)
    b.close(); a.close()
    assert not os.path.exists("/dev/shm" + name)


@pytest.mark.parametrize("mode", ["standard", "paper"])
def test_render_multi_on_one_device_equals_rt_render(gpu, mode):
    sc = gpu.load_scene_from_json(scene_path("quirks"))
    want = staged(gpu, sc, mode)
    got, st = gpu.render_multi(sc, gpu.RenderMode(MODES[mode]), [0])                 # pageable target: library frame + copy
    assert np.array_equal(got, want)
    assert st["primary_rays"] == sc.width * sc.height * (8 if mode == "standard" else 1) and st["kernel_ms"] > 0
    frame = gpu.HostFrame(sc.height, sc.width)
    got2, _ = gpu.render_multi(sc, gpu.RenderMode(MODES[mode]), 1, out=frame)        # host frame: zero copy
    assert np.array_equal(got2, want)
    frame.close()


@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs on one box")
@pytest.mark.parametrize("mode", ["standard", "paper"])
@pytest.mark.parametrize("name", ["quirks", "synthetic_b_small", "snorlax_small"])
def test_render_multi_all_devices_equals_one_device(gpu, name, mode):
    sc = gpu.load_scene_from_json(scene_path(name))
    want = staged(gpu, sc, mode)
    n = min(_gpus(), 8)
    frame = gpu.HostFrame(sc.height, sc.width)
    got, st = gpu.render_multi(sc, gpu.RenderMode(MODES[mode]), n, out=frame)
    assert np.array_equal(got, want), f"{(got != want).any(axis=2).sum()} pixels differ on {n} devices"
    if mode == "standard":
        assert st["primary_rays"] == sc.width * sc.height * 8
    else:       # every rank re-traces the one-row halo of its 16-row blocks (visibility only)
        assert sc.width * sc.height <= st["primary_rays"] <= sc.width * sc.height * (1 + 2 / 16) + 2 * sc.width
    frame.close()


def test_cli_ray_gpus(gpu, tmp_path):
    """`RAY_GPUS=N ray scene.json out.png`: same argv, same PNG as one GPU (device list via the environment)."""
    from PIL import Image
    want = staged(gpu, gpu.load_scene_from_json(scene_path("snorlax_small")), "standard")
    for spec in ["1"] + (["%d" % min(_gpus(), 8), "1,0"] if _gpus() >= 2 else []):
        out = str(tmp_path / ("out_%s.png" % spec.replace(",", "_")))
        r = subprocess.run([gpu.RAY_CLI, scene_path("snorlax_small"), out], capture_output=True, text=True, timeout=300,
                           env=dict(os.environ, RAY_GPUS=spec, RAY_STATS="1"))
        assert r.returncode == 0, r.stderr
        assert np.array_equal(np.asarray(Image.open(out)), want), spec
    r = subprocess.run([gpu.RAY_CLI, scene_path("snorlax_small"), str(tmp_path / "x.png")], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, RAY_GPUS="99"))
    assert r.returncode == 3 and "[error]" in r.stderr
