"""The round-level bundle cull (csrc/rt_cull.cuh) only ever rejects primitives that the exact per-ray test would miss as
well, so the scans' results equal the reference's linear scan (scene.cpp:10-42).  The cull is plain arithmetic, so the
check runs on the host: tests/host/cull_conservative.cpp builds random ray bundles (4 scene scales x 5 origin spreads x 6
direction spreads, window finite or not; as central ray + spreads, as camera rays from a direction box, as shadow rays
from an origin box towards a point / directional light) and spheres near-tangent to one of the bundle's rays, and counts
violations."""
import os
import subprocess

from conftest import ROOT


def test_bundle_cull_is_conservative(tmp_path):
    exe = str(tmp_path / "cull_conservative")
    src = os.path.join(ROOT, "tests", "host", "cull_conservative.cpp")
    subprocess.run(["g++", "-std=c++17", "-O2", "-w", "-o", exe, src], check=True)
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout[-2000:]
    cases, rejected, bad = (int(x) for x in out.stdout.strip().splitlines()[-1].split())
    print(f"{cases} bundle x sphere cases, {rejected} rejected, {bad} violations")
    assert bad == 0
    assert cases > 3_000_000 and rejected > cases // 8, "the cull must actually reject something"
