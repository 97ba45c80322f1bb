"""Multi-GPU host logic on CPU: the tile partition covers every pixel exactly once, and a world_size-2 gloo
run of gather_frame reassembles the frame (the same code path bench.py runs over NCCL)."""
import importlib
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from conftest import ROOT

sharding = importlib.import_module("raytracing-project_b200.sharding")


@pytest.mark.parametrize("W,H", [(13, 7), (320, 240), (1024, 768), (156, 1)])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("mode", [0, 1])
def test_partition_is_exact_cover(W, H, world, mode):
    seen = np.zeros(W * H, int)
    sizes = []
    for r in range(world):
        idx = sharding.owned_pixels(W, H, mode, r, world)
        seen[idx] += 1
        sizes.append(idx.size)
    assert (seen == 1).all()
    if mode == 0 and W * H > 156 * world * 4:
        assert max(sizes) - min(sizes) <= 156          # interleaved tiles balance to within one tile


def test_standard_tiles_follow_the_reference_loop_order():
    W, H = 20, 10
    idx = sharding.owned_pixels(W, H, 0, 0, 1)
    # loop pixel p = y*W + x (y from the BOTTOM row) is stored at row H-1-y (tracer.cpp:297)
    assert idx[0] == (H - 1) * W and idx[W - 1] == (H - 1) * W + W - 1 and idx[W] == (H - 2) * W
    own1 = sharding.owned_pixels(W, H, 0, 1, 2)
    p = np.arange(W * H); p = p[(p // 156) % 2 == 1]
    assert np.array_equal(own1, (H - 1 - p // W) * W + p % W)


WORKER = textwrap.dedent("""
    import importlib, os, sys
    import numpy as np, torch, torch.distributed as dist
    sys.path.insert(0, {root!r})
    sharding = importlib.import_module("raytracing-project_b200.sharding")
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    for (W, H, mode) in [(97, 53, 0), (64, 40, 1)]:
        truth = torch.arange(W * H * 4, dtype=torch.int64).remainder(251).to(torch.uint8).view(H, W, 4)
        local = torch.zeros_like(truth)                      # a rank only fills the pixels it owns
        idx = torch.from_numpy(sharding.owned_pixels(W, H, mode, rank, world))
        local.view(-1, 4)[idx] = truth.view(-1, 4)[idx]
        full = sharding.gather_frame(local, W, H, mode, rank, world)
        if rank == 0:
            assert full is not None and torch.equal(full, truth), (W, H, mode)
        else:
            assert full is None
    dist.barrier(); dist.destroy_process_group()
    open(os.path.join({out!r}, "ok%d" % rank), "w").write("ok")
""")


def test_gather_frame_world2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, out=str(tmp_path)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29631", str(script)],
                       capture_output=True, text=True, timeout=300, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()      # (stdout of the two ranks interleaves)
