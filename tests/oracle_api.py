"""ctypes front end for the CHECKERS under oracle/ (test infrastructure only).

* ``Oracle``      — oracle/_build/liboracle.so, the CPU restatement (rt_oracle.cpp).
* ``ref_driver``  — oracle/_ref/ref_driver, the unmodified reference sources + our driver
                    (built in the authoring container; the binary travels to the GPU box).

Nothing in the product package imports this module.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import tempfile
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
REF_DRIVER = os.path.join(ORACLE_DIR, "_ref", "ref_driver")


def build_oracle() -> None:
    """(Re)build the checkers; `make` is incremental.  _ref is only built where /root/reference exists."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "all"], check=True, stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build_oracle()
        L = C.CDLL(LIB_PATH)
        L.orc_scene_load.restype = C.c_void_p
        L.orc_scene_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.orc_scene_free.argtypes = [C.c_void_p]
        L.orc_add_dir_light.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_scene_info.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.orc_render.restype = C.c_int
        L.orc_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_trace_rays.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_mt_raw.argtypes = [C.c_uint32, C.c_uint64, C.c_int, C.c_void_p]
        L.orc_jitter.argtypes = [C.c_uint64, C.c_int, C.c_void_p]
        _lib = L
    return _lib


class OracleError(RuntimeError):
    pass


class Oracle:
    """One loaded scene in the CPU restatement."""

    def __init__(self, json_text: str):
        err = C.create_string_buffer(1024)
        self._h = lib().orc_scene_load(json_text.encode(), err, 1024)
        if not self._h:
            raise OracleError(err.value.decode())
        info = (C.c_int * 6)()
        lib().orc_scene_info(self._h, info)
        self.width, self.height, self.n_objects, self.n_lights, self.recursion, self.n_materials = list(info)

    @classmethod
    def from_file(cls, path: str) -> "Oracle":
        with open(path) as f:
            return cls(f.read())

    def add_dir_light(self, direction, radiance):
        """scene.dir_lights.push_back({dir, radiance}) (scene.h:10-19,37)."""
        d = np.asarray(direction, np.float64); r = np.asarray(radiance, np.float64)
        lib().orc_add_dir_light(self._h, d.ctypes.data_as(C.c_void_p), r.ctypes.data_as(C.c_void_p))

    def close(self):
        if self._h:
            lib().orc_scene_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, mode: str = "standard", rows=None, want_fb=True, want_ids=False, threads: int = 0):
        """Returns dict(fb=float64[H,W,3], ids=int32[H,W,S], t=float64[H,W,S]); rows=(y0,y1) in LOOP order."""
        W, H = self.width, self.height
        m = 1 if mode == "paper" else 0
        S = 1 if m else 8
        y0, y1 = rows if rows is not None else (0, H)
        fb = np.zeros((H, W, 3), np.float64) if want_fb else None
        ids = np.full((H, W, S), -1, np.int32) if want_ids else None
        ts = np.full((H, W, S), np.inf, np.float64) if want_ids else None
        ptr = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
        threads = threads or min(os.cpu_count() or 1, 16)
        n = max(1, min(threads, y1 - y0))
        edges = np.linspace(y0, y1, n + 1).astype(int)

        def band(k):
            lib().orc_render(self._h, m, int(edges[k]), int(edges[k + 1]), ptr(fb), ptr(ids), ptr(ts))

        if n == 1:
            band(0)
        else:  # bands are independent (the RNG is re-seeded and discarded per band); ctypes drops the GIL
            with ThreadPoolExecutor(n) as ex:
                list(ex.map(band, range(n)))
        return {"fb": fb, "ids": ids, "t": ts}

    def trace_rays(self, rays: np.ndarray) -> np.ndarray:
        rays = np.ascontiguousarray(rays, np.float64).reshape(-1, 6)
        out = np.zeros((rays.shape[0], 16), np.float64)
        lib().orc_trace_rays(self._h, rays.shape[0], rays.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        return out


def mt_raw(seed: int, skip: int, n: int) -> np.ndarray:
    out = np.zeros(n, np.uint32)
    lib().orc_mt_raw(seed, skip, n, out.ctypes.data_as(C.c_void_p))
    return out


def jitter(first_sample: int, n: int) -> np.ndarray:
    out = np.zeros((n, 2), np.float64)
    lib().orc_jitter(first_sample, n, out.ctypes.data_as(C.c_void_p))
    return out


def to_byte(fb: np.ndarray) -> np.ndarray:
    """toByte (core.h:313-316): round-half-away-from-zero of clamp01(v)*255."""
    v = np.clip(np.nan_to_num(fb, nan=0.0), 0.0, 1.0) * 255.0
    return np.floor(v + 0.5).astype(np.uint8)


# ---- the unmodified reference (oracle/_ref) ------------------------------------------------------------

def have_ref() -> bool:
    return os.path.exists(REF_DRIVER) and os.access(REF_DRIVER, os.X_OK)


def _read_fb(path):
    with open(path, "rb") as f:
        W, H = np.frombuffer(f.read(8), np.int32)
        return np.frombuffer(f.read(), np.float64).reshape(H, W, 3).copy()


def _read_ids(path):
    with open(path, "rb") as f:
        W, H, S = np.frombuffer(f.read(12), np.int32)
        n = int(W) * int(H) * int(S)
        ids = np.frombuffer(f.read(4 * n), np.int32).reshape(H, W, S).copy()
        ts = np.frombuffer(f.read(8 * n), np.float64).reshape(H, W, S).copy()
        return ids, ts


def ref_render(scene_path: str, mode: str = "standard", rows=None, want_ids=False, timeout=3600, dir_lights=()):
    """Run the reference build on a scene file; returns dict(fb, ids, t, render_ms).
    dir_lights: [(direction xyz, radiance rgb)] pushed into scene.dir_lights (the JSON format has none)."""
    with tempfile.TemporaryDirectory() as td:
        cmd = [REF_DRIVER, scene_path, "--quiet", "--fb", os.path.join(td, "o.fb")]
        for d, r in dir_lights:
            cmd += ["--dirlight"] + [repr(float(v)) for v in list(d) + list(r)]
        if mode == "paper":
            cmd.append("--paper")
        if rows is not None:
            cmd += ["--rows", str(rows[0]), str(rows[1])]
        if want_ids:
            cmd += ["--ids", os.path.join(td, "o.ids")]
        p = subprocess.run(cmd, check=True, capture_output=True, text=True, timeout=timeout)
        info = json.loads(p.stdout.strip().splitlines()[-1])
        out = {"fb": _read_fb(os.path.join(td, "o.fb")), "ids": None, "t": None, "render_ms": info["render_ms"]}
        if want_ids:
            out["ids"], out["t"] = _read_ids(os.path.join(td, "o.ids"))
        return out


def ref_trace_rays(scene_path: str, rays: np.ndarray, dir_lights=()) -> np.ndarray:
    rays = np.ascontiguousarray(rays, np.float64).reshape(-1, 6)
    with tempfile.TemporaryDirectory() as td:
        fi, fo = os.path.join(td, "in.rays"), os.path.join(td, "out.rays")
        with open(fi, "wb") as f:
            f.write(np.int32(rays.shape[0]).tobytes())
            f.write(rays.tobytes())
        extra = []
        for d, r in dir_lights:
            extra += ["--dirlight"] + [repr(float(v)) for v in list(d) + list(r)]
        subprocess.run([REF_DRIVER, scene_path, "--quiet", "--rays", fi, fo] + extra, check=True, capture_output=True)
        with open(fo, "rb") as f:
            n = int(np.frombuffer(f.read(4), np.int32)[0])
            return np.frombuffer(f.read(), np.float64).reshape(n, 16).copy()
