#!/usr/bin/env python3
"""bench.py — headline benchmark of the per-pixel render loop (BASELINE.json: Mrays/s and ms/frame at
3840x2160 on 1/2/4/8 B200 vs the reference CPU on the host).

    python bench.py --gpus N --steps K --warmup W            # our arm (torchrun launches N>1)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code (oracle/_ref)

A "step" is one frame of the workload scene (default scenes/snorlax_4k.json: snorlax at 3840x2160,
recursion 5, 5 point lights — configs[2], the configuration the metric and the 1,000x target are quoted on).
`value` = rays/s with the scene resident on the device (kernel + pixel exchange, CUDA events, max over ranks);
`e2e`   = the same through the C-ABI call rt_render with HOST buffers: JSON parse + scene compile + upload (H2D) + render,
          the RGB8 frame leaving the GPU(s) inside the kernel's own stores into one pinned host frame (D2H over every
          rank's own PCIe link; rt_b200.h "host-direct output"), wall clock around the call + the ranks' barrier.
Rays = every Scene::intersect-equivalent and Scene::occluded-equivalent query (primary + shadow +
reflection + refraction), counted on the device; the CPU legs count the same queries with the oracle.
"""
from __future__ import annotations

import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
DEFAULT_SCENE = os.path.join(ROOT, "scenes", "snorlax_4k.json")
REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")

# algorithmic FLOPs per counted event (FMA = 2), SURVEY.md §8(d) / DESIGN.md §5
FLOP_PRIM_TEST, FLOP_SHADOW, FLOP_QUERY, FLOP_PIXEL = 17, 80, 30, 12


def traffic_of(workload: str, paper: bool):
    """DRAM bytes per launch of the dominant kernel from this round's `ncu --set full` capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(workload + (":paper" if paper else ""))
    except Exception:
        return None


def hbm_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)["hbm_gbs"]
    except Exception:
        return 6650.0      # B200_PROFILING.md fallback


def flops_of(stats: dict, pixels: int) -> float:
    return (FLOP_PRIM_TEST * stats["prim_tests"] + FLOP_SHADOW * stats["shadow_rays"]
            + FLOP_QUERY * (stats["primary_rays"] + stats["secondary_rays"]) + FLOP_PIXEL * pixels)


# ---------------------------------------------------------------------------------------------------------
# CPU legs: the reference's own code (oracle/_ref/ref_driver), rows y % stride == off of the Standard loop
def _oracle_count(scene_path: str, stride: int, offs: list[int]) -> tuple[int, int]:
    """Rays (closest-hit + occlusion queries) and rows of the sampled rows, counted by the oracle (untimed)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_api as oa
    L = oa.lib()
    L.orc_count_rays.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    o = oa.Oracle.from_file(scene_path)
    rays = rows = 0
    res = {}

    def one(off):
        c = np.zeros(3, np.uint64)
        L.orc_count_rays(o._h, 0, 1 << 30, stride, off, c.ctypes.data_as(C.c_void_p))
        res[off] = c

    th = [threading.Thread(target=one, args=(off,)) for off in offs]
    [t.start() for t in th]; [t.join() for t in th]
    for c in res.values():
        rays += int(c[0] + c[1]); rows += int(c[2])
    return rays, rows


def _ref_run(scene_path: str, stride: int, off: int) -> float:
    """One ref_driver process over rows y % stride == off; returns the ms spent inside the rendered rows."""
    p = subprocess.run([REF_DRIVER, scene_path, "--quiet", "--stride", str(stride), str(off)], capture_output=True, text=True, check=True)
    return float(json.loads(p.stdout.strip().splitlines()[-1])["render_ms"])


def _port_run(scene_path: str, stride: int, off: int) -> float:
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_api as oa
    L = oa.lib()
    L.orc_count_rays.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    o = oa.Oracle.from_file(scene_path)
    c = np.zeros(3, np.uint64)
    t0 = time.perf_counter()
    L.orc_count_rays(o._h, 0, 1 << 30, stride, off, c.ctypes.data_as(C.c_void_p))
    return (time.perf_counter() - t0) * 1e3


def cpu_sample(scene_path: str, height: int, threads: int, target_s: float):
    """Time `threads` concurrent single-threaded runs of the reference over disjoint strided row samples.
    Returns dict(value Mrays/s, wall_ms, rays, rows, kind, stride)."""
    have_ref = os.path.exists(REF_DRIVER)
    run = _ref_run if have_ref else _port_run
    # probe: 2 rows, one thread, to size the sample
    probe_stride = max(1, height // 2)
    t_probe = run(scene_path, probe_stride, probe_stride // 2) / 1e3
    rows_probe = len([y for y in range(height) if y % probe_stride == probe_stride // 2])
    per_row = max(t_probe / max(rows_probe, 1), 1e-4)
    rows_per_thread = max(1, int(target_s / per_row))
    stride = max(threads, height // rows_per_thread)
    stride = max(stride, 1)
    offs = [int(k * stride / threads) for k in range(threads)]
    ms = [0.0] * threads

    def one(k):
        ms[k] = run(scene_path, stride, offs[k])

    t0 = time.perf_counter()
    th = [threading.Thread(target=one, args=(k,)) for k in range(threads)]
    [t.start() for t in th]; [t.join() for t in th]
    wall = (time.perf_counter() - t0) * 1e3
    rays, rows = _oracle_count(scene_path, stride, offs)
    busy = max(ms)                                  # the slowest worker's time inside its rows
    return {"value": rays / (busy * 1e-3) / 1e6, "busy_ms": busy, "wall_ms": wall, "rays": rays, "rows": rows,
            "kind": "reference" if have_ref else "port", "stride": stride, "threads": threads}


def cpu_paper_frame(scene_path: str, width: int, height: int, shadow_rays: int | None):
    """Paper mode has no RNG to discard, so the CPU leg renders the WHOLE frame with the reference (one thread, as the
    reference runs).  Its queries: per pixel one closest hit + the shadow scans of the centre ray (trace_paper + shade,
    tracer.cpp:111-120,262-266) and the centre + four neighbour re-traces of get_edge_strength (tracer.cpp:133-178)."""
    have_ref = os.path.exists(REF_DRIVER)
    if have_ref:
        p = subprocess.run([REF_DRIVER, scene_path, "--quiet", "--paper"], capture_output=True, text=True, check=True)
        ms = float(json.loads(p.stdout.strip().splitlines()[-1])["render_ms"])
    else:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_api as oa
        o = oa.Oracle.from_file(scene_path)
        t0 = time.perf_counter(); o.render("paper", threads=1); ms = (time.perf_counter() - t0) * 1e3
    closest = 2 * width * height + 4 * width * height - 2 * width - 2 * height      # neighbours outside the frame are skipped
    if shadow_rays is None:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_api as oa
        L = oa.lib(); L.orc_count_rays.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        shadow_rays = 0      # the oracle's counter walks the Standard loop only; the shadow share is reported by our arm
    rays = closest + shadow_rays
    return {"value": rays / (ms * 1e-3) / 1e6, "busy_ms": ms, "rays": rays, "kind": "reference" if have_ref else "port"}


L2_NOTE = "flushed between timed steps (256 MB memset)"


def workload_config(scene_path: str, paper: bool, world: int, exchange: str | None = None) -> dict:
    """`config` of the JSON line, derived from the scene file alone so that BOTH arms print the same object for the same
    workload (the reference arm is "on your arm's config": the GPU-only keys describe the arm it is compared with)."""
    with open(scene_path) as f:
        sc = json.load(f)
    scr = sc["screen"]
    W = max(1, round(scr["dimensions"][0] * scr.get("dpi", 72))); H = max(1, round(scr["dimensions"][1] * scr.get("dpi", 72)))
    prims = ("sphere", "halfSpace", "pokeball")

    def leaves(node, inside):                     # primitives below a CSG node (the scene compiler's leaf count)
        n = 0
        for k, v in node.items():
            if k in prims:
                n += 1 if inside else 0
            elif k in ("union", "intersection", "difference"):
                n += sum(leaves(c, True) for c in v)
            elif k == "csg":
                n += leaves(v["left"], True) + leaves(v["right"], True)
            elif k in ("translation", "scaling", "rotation"):
                n += leaves(v["subject"], inside)
        return n
    if exchange is None:
        exchange = "none (1 GPU)" if world == 1 else "peer stores into rank 0's frame over NVLink + 4-byte all-reduce"
    return {"workload": os.path.basename(scene_path), "width": W, "height": H, "spp": 1 if paper else 8,
            "mode": "paper" if paper else "standard", "recursion": int(sc.get("medium", {}).get("recursion", 5)),
            "lights": len(sc.get("sources", [])), "objects": len(sc.get("objects", [])),
            "csg_leaves": sum(leaves(o, False) for o in sc.get("objects", [])),
            "parallelism": f"row-tile interleave x{world}", "exchange": exchange, "l2": L2_NOTE}


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    with open(args.scene) as f:
        scr = json.load(f)["screen"]
    W = max(1, round(scr["dimensions"][0] * scr.get("dpi", 72))); H = max(1, round(scr["dimensions"][1] * scr.get("dpi", 72)))
    threads = os.cpu_count() or 1
    if args.paper:
        # whole frames, one per step, one thread (the paper loop has no sequential RNG, but the reference is single-threaded
        # and a frame takes seconds: no sampling needed)
        vals, ms = [], []
        for i in range(min(args.warmup, 1) + min(args.steps, 3)):
            c = cpu_paper_frame(args.scene, W, H, None)
            if i >= min(args.warmup, 1):
                vals.append(c["value"]); ms.append(c["busy_ms"])
        v = float(np.mean(vals))
        print(json.dumps({
            "impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": len(vals),
            "warmup": min(args.warmup, 1), "ms_per_step": float(np.mean(ms)), "ms_per_frame": float(np.mean(ms)), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.scene, True, args.gpus),
            "reference_step": "one whole paper-mode frame, 1 thread (closest-hit queries only: 6 per pixel, tracer.cpp:133-178,262)",
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": 1, "kind": c["kind"], "sample": f"whole {W}x{H} paper frame"},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    target = max(2.0, min(10.0, 150.0 / max(1, args.steps + args.warmup)))
    vals, step_ms, last = [], [], None
    for i in range(args.warmup + args.steps):
        last = cpu_sample(args.scene, H, threads, target)
        if i >= args.warmup:
            vals.append(last["value"]); step_ms.append(last["busy_ms"])
    v = float(np.mean(vals))
    frame_rays_est = last["rays"] / max(last["rows"], 1) * H
    sample = f"{last['rows']} rows (y % {last['stride']} in {threads} offsets) of {W}x{H}, one single-threaded process per core"
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": float(np.mean(step_ms)), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.scene, False, args.gpus),
        "reference_step": "bounded row sample of one frame (the reference cannot render a sub-frame faster than its RNG stream: rows "
                          "are sampled with rng.discard, oracle/ref_driver.cpp), one single-threaded process per host core",
        "ms_per_frame_extrapolated": frame_rays_est / (v * 1e6) * 1e3,
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": last["kind"], "sample": sample},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv"); os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.proc:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def run_ours(args) -> None:
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rt = importlib.import_module("raytracing-project_b200")
    sharding = importlib.import_module("raytracing-project_b200.sharding")
    rt.init(local)
    scene = rt.load_scene_from_json(args.scene)
    W, H = scene.width, scene.height
    mode = rt.RenderMode.Paper if args.paper else rt.RenderMode.Standard
    tracer = rt.Tracer(scene, mode)
    dev = torch.device("cuda", local)
    frame = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    cache: dict = {}
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: every rank's kernel stores its tiles straight into ONE frame in rank 0's HBM (NVLink peer memory,
    # sharding.PeerFrame); the NCCL pack/gather/scatter path remains for boxes where the mapping is refused.
    peer = None
    if world > 1 and args.exchange == "peer":
        try:
            peer = sharding.PeerFrame(rt, W, H, rank, world, dev)
        except RuntimeError as e:
            if rank == 0:
                print("bench.py: %s; using the NCCL gather" % e, file=sys.stderr)
    target = peer.ptr if peer is not None else frame.data_ptr()
    exchange = "none (1 GPU)" if world == 1 else "peer stores into rank 0's frame over NVLink + 4-byte all-reduce" if peer is not None else "NCCL gather of owned pixels"

    def finish():
        if peer is not None:
            peer.complete()
            return peer.tensor()
        return sharding.gather_frame(frame, W, H, int(mode), rank, world, index_cache=cache)

    def step():
        tracer.render_device(target, stream=stream, rank=rank, world=world)
        return finish()

    # rays per frame (deterministic): one untimed render with the device counters read back
    st = dict(tracer.render_device(target, stream=stream, rank=rank, world=world, want_stats=True))
    keys = ("primary_rays", "shadow_rays", "secondary_rays", "prim_tests")
    cnt = torch.tensor([st[k] for k in keys], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(cnt)
    tot = dict(zip(keys, [int(v) for v in cnt.tolist()]))
    rays_per_frame = tot["primary_rays"] + tot["shadow_rays"] + tot["secondary_rays"]
    # Algorithmic work of the frame = the primitive tests of the reference's linear scans (scene.cpp:10-42), counted by the
    # same kernels with the round-level culling switched off (one untimed frame); what the culled kernels execute is less.
    rt.debug_option("cull", 0)
    scene_lin = rt.load_scene_from_json(args.scene)
    st_lin = dict(rt.Tracer(scene_lin, mode).render_device(target, stream=stream, rank=rank, world=world, want_stats=True))
    rt.debug_option("cull", -1)
    torch.cuda.synchronize(); scene_lin.close()
    lin = torch.tensor([st_lin["prim_tests"]], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(lin)
    tests_linear = int(lin.item())
    peak_tflops = rt.measure_fp32_peak()

    for _ in range(max(args.warmup, 3)):
        step()
    # ---- device-resident timing: K steps, L2 flushed between steps, CUDA events on the launching stream ----
    sampler = ClockSampler(local); 
    if rank == 0:
        sampler.start()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for a, k, b in ev:
        flush.zero_()
        a.record()
        tracer.render_device(target, stream=stream, rank=rank, world=world)
        k.record()
        finish()
        b.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else {}
    step_ms = [a.elapsed_time(b) for a, k, b in ev]
    kern_ms = [a.elapsed_time(k) for a, k, b in ev]
    t = torch.tensor([sum(step_ms), sum(kern_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t[0]) / args.steps
    kernel_ms = float(t[1]) / args.steps

    # ---- end to end through the C ABI with host buffers: parse + compile + upload (H2D) + render + the frame in HOST memory ----
    # The host frame is pinned, device-mapped memory (N = 1: a pinned torch tensor; N > 1: ONE POSIX-shm frame every rank
    # maps, rt_host_frame_*): rt_render's kernel stores the packed RGB8 pixels straight into it, every rank over its own
    # PCIe link, and the step ends with the ranks' barrier.  Fallback (no /dev/shm): peer frame + one D2H copy on rank 0.
    shared = None; e2e_path = "rt_render -> pinned host tensor (host-direct stores)"
    if world > 1:
        name = "/rt_b200_%s_%d" % (os.environ.get("MASTER_PORT", "0"), os.getppid())
        ok = torch.ones(1, dtype=torch.int32, device=dev)
        try:
            if rank == 0:
                shared = rt.HostFrame(H, W, name=name, create=True)
        except Exception as e:      # noqa: BLE001
            print("bench.py: shared host frame unavailable (%s)" % e, file=sys.stderr); ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) and rank != 0:
            try:
                shared = rt.HostFrame(H, W, name=name, create=False)
            except Exception as e:  # noqa: BLE001
                print("bench.py: rank %d cannot map the shared host frame (%s)" % (rank, e), file=sys.stderr); ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if not int(ok.item()):
            if shared is not None:
                shared.close()
            shared = None
        e2e_path = ("rt_render(rank, world) on every rank -> ONE shared pinned host frame (host-direct stores) + barrier" if shared is not None
                    else "rt_render_device -> peer frame on rank 0 -> one D2H copy")
    host = None
    if shared is None:
        host = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() if world == 1 else torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
    d = scene.desc
    h2d = (d.n_objects + d.n_leaves) * 32 + d.n_pokeballs * 48 + d.n_affines * 48 + d.n_ops * 4 + d.n_materials * 48 + d.n_lights * 32 + 256
    def e2e_step():
        s2 = rt.load_scene_from_json(args.scene)                         # parse + compile on the host, upload below
        if world == 1 or shared is not None:
            dst = host.data_ptr() if shared is None else shared.ptr
            rt._check(rt.lib().rt_render(s2.device_handle(), int(mode), rank, world, C.c_void_p(dst), None, None, None))
            if world > 1:
                dist.barrier()
        else:
            t2 = rt.Tracer(s2, mode)
            t2.render_device(target, stream=stream, rank=rank, world=world)
            full = finish()
            if rank == 0:
                host.copy_(full, non_blocking=True)
            torch.cuda.synchronize()
        s2.close()
    for _ in range(2):
        e2e_step()
    barrier(); t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier(); e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    te = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_ms = float(te[0])
    d2h_bytes = W * H * (3 if (world == 1 or shared is not None) else 4)

    # ---- N > 1: the peer-assembled frame must equal the frame one GPU renders alone, byte for byte (untimed) ----
    frame_equal = None
    if world > 1:
        assembled = step()
        assembled = assembled.clone() if assembled is not None else None
        torch.cuda.synchronize()
        if rank == 0:
            alone = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
            tracer.render_device(alone.data_ptr(), stream=stream, rank=0, world=1)
            torch.cuda.synchronize()
            frame_equal = bool(torch.equal(assembled.reshape(H, W, 4), alone))
            if shared is not None:      # ... and so must the frame the e2e leg assembled in host memory
                frame_equal = frame_equal and bool(np.array_equal(shared.array, alone[..., :3].cpu().numpy()))
        barrier()
    elif host is not None:
        alone = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
        tracer.render_device(alone.data_ptr(), stream=stream, rank=0, world=1)
        torch.cuda.synchronize()
        frame_equal = bool(np.array_equal(host.numpy(), alone[..., :3].cpu().numpy()))     # host-direct frame == device frame

    if rank == 0:
        flops_exec = flops_of(tot, W * H)
        flops = flops_of(dict(tot, prim_tests=max(tests_linear, tot["prim_tests"])), W * H)
        achieved = flops / (kernel_ms * 1e-3) / 1e12 / world          # per GPU
        line = {
            "metric": "Mrays/s", "value": rays_per_frame / (ms_per_step * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "ms_per_frame": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.scene, args.paper, world, exchange),
            "rays_per_frame": rays_per_frame, "rays": tot, "kernel_ms": kernel_ms,
            "e2e": {"value": rays_per_frame / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": e2e_ms,
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h_bytes), "path": e2e_path},
            "gpu_launches": args.steps * (2 if args.paper else 1) * world,
            # north_star: this path is FP32 CUDA-core work, reported against the FP32 roofline.  MEASURED_PEAKS.json holds
            # only HBM and bf16-tensor peaks, so the denominator is measured live on this GPU (FFMA micro-benchmark).
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                         "traffic": traffic_of(os.path.basename(args.scene), args.paper),
                         "peak_source": "measured live (dependent-free FFMA micro-benchmark, rt_measure_fp32_peak)",
                         "flops_per_frame": flops,
                         "flops_note": "algorithmic = 17 x the primitive tests of the reference's linear scans (same kernels with "
                                       "round-level culling off, one untimed frame) + 80/shadow ray + 30/closest hit + 12/pixel; "
                                       "flops_executed_per_frame counts the tests the culled kernels really ran",
                         "prim_tests_linear_scan": tests_linear, "flops_executed_per_frame": flops_exec,
                         "achieved_executed": flops_exec / (kernel_ms * 1e-3) / 1e12 / world,
                         "framebuffer": {"bytes_per_frame": int(W * H * 4), "achieved_gbs": W * H * 4 / (kernel_ms * 1e-3) / 1e9 / world,
                                         "hbm_peak_gbs": hbm_peak_gbs(), "note": "RGBA8 stores only; the frame is compute-bound"}},
            "clocks": clocks,
        }
        if frame_equal is not None:
            line["frame_equals_single_gpu" if world > 1 else "host_frame_equals_device_frame"] = frame_equal
        if world == 1 and not args.no_cpu:
            try:
                if args.paper:
                    c = cpu_paper_frame(args.scene, W, H, tot["shadow_rays"])
                    line["cpu_baseline"] = {"value": c["value"], "unit": "Mrays/s", "cores": 1, "kind": c["kind"],
                                            "sample": f"whole {W}x{H} paper frame, 1 thread; the reference re-traces 5 extra closest hits per pixel "
                                                      f"(tracer.cpp:133-178), counted as its rays", "ms_per_frame": c["busy_ms"]}
                else:
                    c = cpu_sample(args.scene, H, 1, args.cpu_seconds)
                    line["cpu_baseline"] = {"value": c["value"], "unit": "Mrays/s", "cores": 1, "kind": c["kind"],
                                            "sample": f"{c['rows']} rows (y % {c['stride']} == 0) of {W}x{H}, 1 thread, {c['busy_ms']:.0f} ms inside the rows",
                                            "ms_per_frame_extrapolated": rays_per_frame / (c["value"] * 1e6) * 1e3}
            except Exception as e:  # the baseline is a report, never a reason to lose the GPU number
                line["cpu_baseline"] = {"error": str(e)}
        print(json.dumps(line))
    if shared is not None:
        barrier()
        shared.close()
    if peer is not None:
        peer.close()
    if world > 1:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default=DEFAULT_SCENE)
    ap.add_argument("--paper", action="store_true")
    ap.add_argument("--exchange", default="peer", choices=["peer", "gather"], help="N>1: how pixels reach rank 0")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
