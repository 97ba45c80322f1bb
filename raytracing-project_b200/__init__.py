"""raytracing-project_b200 — host-side mirror (Python) of the reference's render interface over the
C ABI in include/rt_b200.h.

The reference's public surface for the hot path is ``jsonio::load_scene_from_json`` +
``Tracer{scene,camera,width,height,mode}.render(framebuffer)`` (raytracer/src/json_loader.h:27-36,
raytracer/src/tracer.h:18-35).  The same names are kept here:

    scene = load_scene_from_json("examples/pokeballs.json")      # or load_scene_from_json_text(text)
    fb = Tracer(scene, mode=RenderMode.Standard).render()         # float32 [H, W, 3], row 0 = top

The product's own host language is C++17 (csrc/ray_main.cpp is the CLI); this module exists so the
parity tests and bench.py can call the very same shared library.  There is NO CPU fallback: if
``librt_b200.so`` is missing or no CUDA device is present every render call raises.
"""
from __future__ import annotations

import ctypes as C
import enum
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librt_b200.so")
RAY_CLI = os.path.join(_HERE, "ray")
TILE_PIXELS = 156
PAPER_ROWS = 16
NOALLOC_MAX_PIXELS = 1280 * 1024     # RT_NOALLOC_MAX_PIXELS


class RtError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"rt_b200 error {code}: {message}")
        self.code = code
        self.message = message


class RenderMode(enum.IntEnum):      # tracer.h:8-11
    Standard = 0
    Paper = 1


class RtStats(C.Structure):
    _fields_ = [("kernel_ms", C.c_double), ("total_ms", C.c_double),
                ("primary_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("secondary_rays", C.c_uint64),
                ("prim_tests", C.c_uint64),
                ("width", C.c_int32), ("height", C.c_int32), ("samples_per_pixel", C.c_int32), ("n_tiles", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RtPrim(C.Structure):
    _fields_ = [("geom", C.c_float * 4), ("kind", C.c_int32), ("flags", C.c_int32), ("mat", C.c_int32),
                ("aux", C.c_int32), ("xf", C.c_int32), ("pad", C.c_int32 * 3)]


class RtMaterial(C.Structure):
    _fields_ = [("albedo", C.c_float * 3), ("ks", C.c_float), ("ambient", C.c_float * 3), ("kr", C.c_float),
                ("kt", C.c_float), ("shininess", C.c_float), ("ior", C.c_float), ("kd", C.c_float)]


class RtLight(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("pad0", C.c_float), ("intensity", C.c_float * 3), ("pad1", C.c_float)]


class RtDirLight(C.Structure):
    _fields_ = [("to_light", C.c_float * 3), ("pad0", C.c_float), ("radiance", C.c_float * 3), ("pad1", C.c_float)]


class RtPokeball(C.Structure):
    _fields_ = [("btn_dir", C.c_float * 3), ("cos_outer", C.c_float), ("up", C.c_float * 3), ("cos_inner", C.c_float),
                ("belt_half", C.c_float), ("pad", C.c_float * 3)]


class RtAffine(C.Structure):
    _fields_ = [("m", (C.c_float * 4) * 3)]


class RtSceneDesc(C.Structure):
    _fields_ = [("eye", C.c_double * 3), ("screen_p", C.c_double * 3), ("lx", C.c_double), ("ly", C.c_double),
                ("width", C.c_int32), ("height", C.c_int32),
                ("background", C.c_float * 3), ("medium_index", C.c_float),
                ("recursion", C.c_int32),
                ("n_objects", C.c_int32), ("n_leaves", C.c_int32), ("n_pokeballs", C.c_int32), ("n_affines", C.c_int32),
                ("n_ops", C.c_int32), ("n_materials", C.c_int32), ("n_lights", C.c_int32), ("n_fallbacks", C.c_int32),
                ("objects", C.POINTER(RtPrim)), ("leaves", C.POINTER(RtPrim)), ("pokeballs", C.POINTER(RtPokeball)),
                ("affines", C.POINTER(RtAffine)), ("ops", C.POINTER(C.c_uint32)), ("materials", C.POINTER(RtMaterial)),
                ("lights", C.POINTER(RtLight)), ("fallbacks", C.POINTER(C.c_float)),
                ("n_dir_lights", C.c_int32), ("pad0", C.c_int32), ("dir_lights", C.POINTER(RtDirLight))]


# every symbol include/rt_b200.h declares (tests/test_abi.py checks the list against the header)
ABI_SYMBOLS = [
    "rt_init", "rt_init_devices", "rt_shutdown", "rt_last_error", "rt_version",
    "rt_host_frame_alloc", "rt_host_frame_open", "rt_host_frame_free", "rt_render_multi",
    "rt_compile_json_text", "rt_compile_json_file", "rt_compiled_desc", "rt_compiled_free", "rt_compiled_add_dir_light",
    "rt_scene_upload", "rt_scene_free", "rt_scene_size",
    "rt_render", "rt_render_device", "rt_render_paper_device_noalloc", "rt_trace_rays", "rt_debug_jitter", "rt_debug_mt_block_host", "rt_debug_mt_jump_host", "rt_debug_device_layout", "rt_debug_mask_table", "rt_debug_set_option",
    "rt_measure_fp32_peak",
    "rt_write_png_rgb8",
    "rt_peer_frame_create", "rt_peer_frame_open", "rt_peer_frame_close",
]

_lib = None


def build(verbose: bool = False) -> None:
    """Compile librt_b200.so and the `ray` CLI in-tree (nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo)."""
    subprocess.run(["make", "-C", _HERE] + ([] if verbose else ["-s"]), check=True)


def lib():
    """The loaded C-ABI library.  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtError(-6, f"{LIB_PATH} is missing: run `make -C {_HERE}` (there is no fallback path)")
        L = C.CDLL(LIB_PATH)
        L.rt_last_error.restype = C.c_char_p
        L.rt_version.restype = C.c_char_p
        L.rt_init.argtypes = [C.c_int]
        L.rt_compile_json_text.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.rt_compile_json_file.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.rt_compiled_desc.restype = C.POINTER(RtSceneDesc)
        L.rt_compiled_desc.argtypes = [C.c_void_p]
        L.rt_compiled_free.argtypes = [C.c_void_p]
        L.rt_compiled_add_dir_light.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.rt_scene_upload.argtypes = [C.POINTER(RtSceneDesc), C.POINTER(C.c_void_p)]
        L.rt_scene_free.argtypes = [C.c_void_p]
        L.rt_scene_size.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.rt_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(RtStats)]
        L.rt_render_device.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.POINTER(RtStats)]
        L.rt_render_paper_device_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.rt_trace_rays.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.rt_debug_jitter.argtypes = [C.c_uint64, C.c_int, C.c_void_p]
        L.rt_debug_mt_block_host.argtypes = [C.c_uint32, C.c_uint64, C.c_void_p]
        L.rt_debug_mt_jump_host.argtypes = [C.c_uint32, C.c_int, C.c_void_p]
        L.rt_measure_fp32_peak.argtypes = [C.POINTER(C.c_double)]
        L.rt_debug_device_layout.argtypes = [C.POINTER(RtSceneDesc), C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.POINTER(C.c_int), C.c_void_p]
        L.rt_write_png_rgb8.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
        L.rt_init_devices.argtypes = [C.c_void_p, C.c_int]
        L.rt_host_frame_alloc.argtypes = [C.c_size_t, C.c_char_p, C.POINTER(C.c_void_p)]
        L.rt_host_frame_open.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.rt_host_frame_free.argtypes = [C.c_void_p]
        L.rt_render_multi.argtypes = [C.POINTER(RtSceneDesc), C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(RtStats)]
        L.rt_peer_frame_create.argtypes = [C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]
        L.rt_peer_frame_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.rt_peer_frame_close.argtypes = [C.c_void_p, C.c_int]
        _lib = L
    return _lib


def _check(rc: int) -> None:
    if rc != 0:
        raise RtError(rc, lib().rt_last_error().decode(errors="replace"))


def init(device: int = -1) -> None:
    _check(lib().rt_init(device))


def shutdown() -> None:
    lib().rt_shutdown()


class Scene:
    """A compiled scene (host) and, once a render is asked for, its device copy.

    Mirrors what jsonio::load_scene_from_json fills (Scene + Camera, json_loader.cpp:458-504)."""

    def __init__(self, json_text: str | None = None, path: str | None = None):
        self._compiled = C.c_void_p()
        self._device = C.c_void_p()
        if path is not None:
            _check(lib().rt_compile_json_file(os.fsencode(path), C.byref(self._compiled)))
        else:
            _check(lib().rt_compile_json_text((json_text or "").encode(), C.byref(self._compiled)))
        self.desc = lib().rt_compiled_desc(self._compiled).contents
        self.width, self.height = int(self.desc.width), int(self.desc.height)

    # -- flat description, for the host-logic tests ---------------------------------------------
    def objects(self):
        return [RtPrim.from_buffer_copy(self.desc.objects[i]) for i in range(self.desc.n_objects)]

    def leaves(self):
        return [RtPrim.from_buffer_copy(self.desc.leaves[i]) for i in range(self.desc.n_leaves)]

    def ops(self):
        return [int(self.desc.ops[i]) for i in range(self.desc.n_ops)]

    def materials(self):
        return [RtMaterial.from_buffer_copy(self.desc.materials[i]) for i in range(self.desc.n_materials)]

    def device_layout(self):
        """(order, runs[n,4], clusters[m,4]) of the device layout (host-only; see rt_debug_device_layout)."""
        n = int(self.desc.n_objects)
        order = np.zeros(max(n, 1), np.int32); runs = np.zeros((max(n, 1), 4), np.int32)
        clusters = np.zeros((n + n // 16 + 2, 4), np.float32); nr, nc = C.c_int(), C.c_int()
        _check(lib().rt_debug_device_layout(C.byref(self.desc), order.ctypes.data_as(C.c_void_p), C.byref(nr),
                                            runs.ctypes.data_as(C.c_void_p), C.byref(nc), clusters.ctypes.data_as(C.c_void_p)))
        return order[:n], runs[: nr.value], clusters[: nc.value]

    def mask_table(self):
        """(masks[n,4], scale): the table of the device's conservative candidate scan (host-only; see rt_debug_mask_table)."""
        n, scale = C.c_int(), C.c_float()
        _check(lib().rt_debug_mask_table(C.byref(self.desc), C.byref(n), None, C.byref(scale)))
        masks = np.zeros((max(n.value, 1), 4), np.float32)
        _check(lib().rt_debug_mask_table(C.byref(self.desc), C.byref(n), masks.ctypes.data_as(C.c_void_p), C.byref(scale)))
        return masks[: n.value], float(scale.value)

    def add_dir_light(self, direction, radiance) -> None:
        """scene.dir_lights.push_back({dir, radiance}) (scene.h:10-19,37); before the first render."""
        if self._device:
            raise RtError(-6, "add_dir_light: the scene is already on the device")
        d = np.asarray(direction, np.float64); r = np.asarray(radiance, np.float64)
        _check(lib().rt_compiled_add_dir_light(self._compiled, d.ctypes.data_as(C.c_void_p), r.ctypes.data_as(C.c_void_p)))

    def device_handle(self):
        if not self._device:
            _check(lib().rt_scene_upload(C.byref(self.desc), C.byref(self._device)))
        return self._device

    def close(self):
        if getattr(self, "_device", None):
            lib().rt_scene_free(self._device)
            self._device = C.c_void_p()
        if getattr(self, "_compiled", None):
            lib().rt_compiled_free(self._compiled)
            self._compiled = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def load_scene_from_json(filename: str) -> Scene:          # json_loader.h:27
    return Scene(path=filename)


def load_scene_from_json_text(json_text: str) -> Scene:    # json_loader.h:36
    return Scene(json_text=json_text)


class Tracer:
    """Tracer (tracer.h:18-35): ``render()`` returns the linear framebuffer, row 0 = top."""

    def __init__(self, scene: Scene, mode: RenderMode = RenderMode.Standard):
        self.scene = scene
        self.mode = RenderMode(mode)
        self.width, self.height = scene.width, scene.height
        self.last_stats: dict | None = None

    def render(self, want_rgb8: bool = False, want_ids: bool = False, rank: int = 0, world: int = 1):
        """Returns fb float32[H,W,3]; with want_rgb8/want_ids a dict(fb=, rgb8=uint8[H,W,3], ids=int32[H,W,S])."""
        W, H = self.width, self.height
        S = 1 if self.mode == RenderMode.Paper else 8
        fb = np.zeros((H, W, 3), np.float32)
        rgb8 = np.zeros((H, W, 3), np.uint8) if want_rgb8 else None
        ids = np.full((H, W, S), -1, np.int32) if want_ids else None
        st = RtStats()
        ptr = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
        _check(lib().rt_render(self.scene.device_handle(), int(self.mode), rank, world, ptr(rgb8), ptr(fb), ptr(ids), C.byref(st)))
        self.last_stats = st.as_dict()
        if want_rgb8 or want_ids:
            return {"fb": fb, "rgb8": rgb8, "ids": ids}
        return fb

    def render_device(self, rgba8_ptr: int, rgbf_ptr: int = 0, ids_ptr: int = 0, stream: int | None = None,
                      rank: int = 0, world: int = 1, want_stats: bool = False):
        """Render into caller-owned DEVICE buffers (e.g. torch tensors' data_ptr()); asynchronous unless want_stats.

        stream: a cudaStream_t as an integer (torch.cuda.current_stream().cuda_stream); 0 means the legacy default
        stream (passed to the library as cudaStreamLegacy); None means the library's own stream."""
        if stream is not None and stream == 0:
            stream = 1          # cudaStreamLegacy: NULL is "library stream" in the C ABI
        st = RtStats()
        _check(lib().rt_render_device(self.scene.device_handle(), int(self.mode), rank, world,
                                      C.c_void_p(rgba8_ptr), C.c_void_p(rgbf_ptr or None), C.c_void_p(ids_ptr or None),
                                      C.c_void_p(stream or None), C.byref(st) if want_stats else None))
        if want_stats:
            self.last_stats = st.as_dict()
        return self.last_stats

    def render_paper_device_noalloc(self, rgba8_ptr: int, stream: int | None = None, rank: int = 0, world: int = 1) -> None:
        """Paper frame into a caller-owned DEVICE buffer through the entry point that allocates nothing
        (rt_render_paper_device_noalloc: G-buffer in a static device array, at most NOALLOC_MAX_PIXELS pixels)."""
        if stream is not None and stream == 0:
            stream = 1          # cudaStreamLegacy, as in render_device
        _check(lib().rt_render_paper_device_noalloc(self.scene.device_handle(), rank, world, C.c_void_p(rgba8_ptr), C.c_void_p(stream or None)))

    def trace_rays(self, rays: np.ndarray) -> np.ndarray:
        """Known-answer entry point: [n,6] float64 rays -> [n,16] float64 records (see rt_b200.h)."""
        rays = np.ascontiguousarray(rays, np.float64).reshape(-1, 6)
        out = np.zeros((rays.shape[0], 16), np.float64)
        _check(lib().rt_trace_rays(self.scene.device_handle(), rays.shape[0], rays.ctypes.data_as(C.c_void_p),
                                   out.ctypes.data_as(C.c_void_p)))
        return out


def debug_jitter(first_sample: int, n: int) -> np.ndarray:
    out = np.zeros((n, 2), np.float64)
    _check(lib().rt_debug_jitter(first_sample, n, out.ctypes.data_as(C.c_void_p)))
    return out


def mt_block_host(seed: int, block: int) -> np.ndarray:
    out = np.zeros(624, np.uint32)
    _check(lib().rt_debug_mt_block_host(seed, block, out.ctypes.data_as(C.c_void_p)))
    return out


def mt_jump_host(seed: int, log2_twists: int) -> np.ndarray:
    """Outputs of block 2**log2_twists reached by ONE polynomial jump on the host (see rt_debug_mt_jump_host)."""
    out = np.zeros(624, np.uint32)
    _check(lib().rt_debug_mt_jump_host(seed, log2_twists, out.ctypes.data_as(C.c_void_p)))
    return out


DEBUG_OPTIONS = ("flat", "cull", "wavefront", "global_scene", "mt_sequential", "grid", "grid_cells")


def debug_option(name: str, value: int = -1) -> None:
    """Test hook (rt_debug_set_option): force a kernel / data path for scenes uploaded afterwards; -1 restores the default."""
    lib().rt_debug_set_option.argtypes = [C.c_char_p, C.c_int]
    _check(lib().rt_debug_set_option(name.encode(), int(value)))


def measure_fp32_peak() -> float:
    v = C.c_double()
    _check(lib().rt_measure_fp32_peak(C.byref(v)))
    return float(v.value)


class HostFrame:
    """Pinned host memory mapped into the device's address space (rt_host_frame_*): rt_render / rt_render_multi store the
    packed RGB8 pixels straight into it.  `name` (e.g. "/rt_frame_1234"): backed by POSIX shared memory, so that the ranks
    of a multi-process job fill ONE frame (rank 0 creates, the others pass create=False)."""

    def __init__(self, height: int, width: int, name: str | None = None, create: bool = True):
        self.shape = (height, width, 3)
        self.nbytes = height * width * 3
        ptr = C.c_void_p()
        if create:
            _check(lib().rt_host_frame_alloc(self.nbytes, name.encode() if name else None, C.byref(ptr)))
        else:
            _check(lib().rt_host_frame_open(name.encode(), self.nbytes, C.byref(ptr)))
        self.ptr = int(ptr.value)
        self.array = np.ctypeslib.as_array((C.c_uint8 * self.nbytes).from_address(self.ptr)).reshape(self.shape)

    def close(self) -> None:
        if getattr(self, "ptr", 0):
            self.array = None
            _check(lib().rt_host_frame_free(C.c_void_p(self.ptr)))
            self.ptr = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def render_multi(scene: "Scene", mode: "RenderMode", devices, out=None):
    """Tracer::render on several GPUs from this one process (rt_render_multi).  devices: a count or a list of ordinals;
    out: a HostFrame (zero-copy) or None (a numpy array is returned).  Returns (rgb8[H,W,3], stats dict)."""
    devs = list(range(devices)) if isinstance(devices, int) else list(devices)
    arr = (C.c_int * len(devs))(*devs)
    st = RtStats()
    target = out.array if out is not None else np.zeros((scene.height, scene.width, 3), np.uint8)
    _check(lib().rt_render_multi(C.byref(scene.desc), int(mode), arr, len(devs), target.ctypes.data_as(C.c_void_p), C.byref(st)))
    return target, st.as_dict()


IPC_HANDLE_BYTES = 64


def peer_frame_create(nbytes: int):
    """Zero-filled device allocation that other processes can map (rank 0 of a multi-GPU frame): (ptr, handle bytes)."""
    ptr = C.c_void_p()
    h = C.create_string_buffer(IPC_HANDLE_BYTES)
    _check(lib().rt_peer_frame_create(nbytes, C.byref(ptr), h))
    return int(ptr.value), h.raw


def peer_frame_open(handle: bytes) -> int:
    """Map another rank's peer frame into this process (NVLink peer memory); returns the device pointer."""
    if len(handle) != IPC_HANDLE_BYTES:
        raise ValueError("IPC handle must be %d bytes" % IPC_HANDLE_BYTES)
    ptr = C.c_void_p()
    _check(lib().rt_peer_frame_open(C.create_string_buffer(handle, IPC_HANDLE_BYTES), C.byref(ptr)))
    return int(ptr.value)


def peer_frame_close(ptr: int, owner: bool) -> None:
    _check(lib().rt_peer_frame_close(C.c_void_p(ptr), 1 if owner else 0))


def write_png_rgb8(path: str, rgb8: np.ndarray) -> None:
    rgb8 = np.ascontiguousarray(rgb8, np.uint8)
    H, W = rgb8.shape[:2]
    _check(lib().rt_write_png_rgb8(os.fsencode(path), rgb8.ctypes.data_as(C.c_void_p), W, H))
