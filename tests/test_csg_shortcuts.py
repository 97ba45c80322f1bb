"""csg_combine's shortcuts (one child missed; both hit with well-separated ends) return exactly what the literal
event sweep of CSG::interval (csg.cpp:61-163: tolerance comparator, libstdc++ insertion sort, first inside run) returns.
The interval arithmetic of csrc/rt_csg.cuh is plain C++, so the check runs on the host: tests/host/csg_shortcuts.cpp walks
a grid of adversarial interval ends (ties, origin, inverted, +-inf, NaN) x 3 operations x hit/miss flags, then 3 M random
intervals."""
import os
import subprocess

from conftest import ROOT


def test_shortcuts_equal_the_literal_sweep(tmp_path):
    exe = str(tmp_path / "csg_shortcuts")
    src = os.path.join(ROOT, "tests", "host", "csg_shortcuts.cpp")
    subprocess.run(["g++", "-std=c++17", "-O2", "-w", "-o", exe, src], check=True)
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout[-2000:]
    cases, bad, fast = (int(x) for x in out.stdout.strip().splitlines()[-1].split())
    assert bad == 0
    assert cases > 6_000_000 and fast > 1_000_000, "the random part must exercise the ordered shortcut"
