"""The uniform grid for incoherent rays (csrc/rt_grid.cuh) only ever drops spheres that the exact per-ray test misses as
well, so the scans' results equal the reference's linear scan (scene.cpp:10-42).  Builder and walk are plain arithmetic shared
by host and device, so the check runs on the host: tests/host/grid_conservative.cpp builds random sphere sets (lattices like
the glass scene, random clouds, four scales), walks random segments (origins outside / inside / on a sphere; random,
axis-parallel, planar and near-tangent directions; finite and infinite windows) and counts accepted hits whose bit is missing."""
import os
import subprocess

from conftest import ROOT


def test_grid_walk_is_conservative(tmp_path):
    exe = str(tmp_path / "grid_conservative")
    src = os.path.join(ROOT, "tests", "host", "grid_conservative.cpp")
    subprocess.run(["g++", "-std=c++17", "-O2", "-w", "-o", exe, src], check=True)
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, (out.stdout + out.stderr)[-2000:]
    rays, hits, cells100, bits100, bad = (int(x) for x in out.stdout.strip().splitlines()[-1].split())
    print(f"{rays} segments, {hits} accepted sphere hits, {cells100 / 100} cells and {bits100 / 100} candidate bits per segment, {bad} violations")
    assert bad == 0
    assert rays >= 1_000_000 and hits > rays, "the segments must actually hit spheres"
    assert bits100 < 20 * 100, "the walk must select a small subset of the ~75 spheres"
