"""Host logic of the device layout (no GPU): clustered runs are a permutation inside their run, every cluster's
bounding sphere contains its 16 members with margin, and short / dense runs stay in JSON order."""
import json

import numpy as np
import pytest

from conftest import scene_path


def spheres_of(scene):
    return scene.objects()


@pytest.mark.parametrize("name", ["synthetic_a_small", "synthetic_a_tiny", "synthetic_b_small", "snorlax_small", "pokeballs_small"])
def test_layout_is_a_permutation_within_runs(rt, name):
    sc = rt.load_scene_from_json(scene_path(name))
    order, runs, clusters = sc.device_layout()
    n = sc.desc.n_objects
    assert sorted(order.tolist()) == list(range(n))
    covered = 0
    for kind, first, count, ctab in runs.tolist():
        assert first == covered
        covered += count
        members = order[first:first + count]
        assert sorted(members.tolist()) == list(range(first, first + count)), "a run keeps its own objects"
        if ctab == 0:
            assert members.tolist() == list(range(first, first + count)), "unclustered runs stay in JSON order"
    assert covered == n


def test_cluster_bounds_contain_their_members(rt):
    sc = rt.load_scene_from_json(scene_path("synthetic_a_small"))
    order, runs, clusters = sc.device_layout()
    objs = sc.objects()
    clustered = [r for r in runs.tolist() if r[3] > 0 and r[0] != 3]
    assert len(clustered) == 2 and [r[2] for r in clustered] == [1000, 64]      # 1 000 spheres and 64 pokeballs
    generic = [r for r in runs.tolist() if r[0] == 3]
    assert len(generic) == 1 and generic[0][2] == 32 and generic[0][3] > 0        # 32 CSG programs: one world bound each
    assert len(clusters) == 63 + 4 + 32
    g0 = generic[0]
    for k in range(32):
        prog = objs[g0[1] + k]
        assert prog.kind == 3 and list(clusters[g0[3] - 1 + k]) == pytest.approx(list(prog.geom))
    for kind, first, count, ctab in clustered:
        for c in range((count + 15) // 16):
            cx, cy, cz, R = clusters[ctab - 1 + c]
            for i in range(first + 16 * c, min(first + 16 * c + 16, first + count)):
                g = objs[int(order[i])].geom
                dist = np.sqrt((g[0] - cx) ** 2 + (g[1] - cy) ** 2 + (g[2] - cz) ** 2)
                assert dist + abs(g[3]) <= R * (1 - 1e-3), "bound must contain the member with margin"
    # spatial clustering must actually be tight: mean cluster radius well below the run's extent
    assert float(np.mean(clusters[:63, 3])) < 3.0


def test_dense_or_short_runs_are_not_clustered(rt):
    """Sphere / pokeball runs that are short or dense stay in JSON order without cluster bounds; every generic object
    (CSG program, wrapped / affine primitive) has exactly one world bound (R = inf when nothing is known)."""
    for name in ("synthetic_b_small", "snorlax_small", "penguin_small", "quirks"):
        sc = rt.load_scene_from_json(scene_path(name))
        order, runs, clusters = sc.device_layout()
        assert order.tolist() == list(range(sc.desc.n_objects))
        n_generic = sum(r[2] for r in runs.tolist() if r[0] == 3)
        assert len(clusters) == n_generic
        for kind, first, count, ctab in runs.tolist():
            assert (ctab > 0) == (kind == 3)
    objs = rt.load_scene_from_json(scene_path("snorlax_small")).objects()
    _, runs, clusters = rt.load_scene_from_json(scene_path("snorlax_small")).device_layout()
    g = [r for r in runs.tolist() if r[0] == 3][0]
    assert list(clusters[g[3] - 1]) == pytest.approx(list(objs[g[1]].geom)), "the program's bound is its compiled bounding sphere"


def test_layout_of_an_empty_and_a_non_finite_scene(rt):
    base = {"screen": {"dpi": 10, "dimensions": [1, 1], "position": [0, 0, 0], "observer": [0, 0, 1]}, "objects": []}
    order, runs, clusters = rt.load_scene_from_json_text(json.dumps(base)).device_layout()
    assert len(order) == 0 and len(runs) == 0 and len(clusters) == 0
    base["objects"] = [{"sphere": {"position": [1e39 if i == 3 else float(i), 0, -5], "radius": 0.1, "color": {"diffuse": [1, 1, 1]}}} for i in range(200)]
    order, runs, clusters = rt.load_scene_from_json_text(json.dumps(base)).device_layout()
    assert len(clusters) == 0 and order.tolist() == list(range(200)), "a run with a non-finite centre is left alone"
