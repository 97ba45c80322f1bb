"""Two-GPU parity (skipped on a one-GPU box): tiles rendered by several ranks, assembled either by peer stores into
rank 0's frame or by the NCCL gather, give the frame one GPU renders alone (tracer.cpp:288-299 is one loop)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT


def _gpus() -> int:
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs on one box")
def test_peer_frame_and_gather_match_single_gpu():
    n = min(_gpus(), 8)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", "29641", os.path.join(ROOT, "scripts", "peer_frame_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("peer==alone True") == 3 and "False" not in r.stdout
