"""Known answers the reference's own Catch2 tests pin for the hot path (raytracer/tests/test_geometry.cpp,
test_csg.cpp, test_camera.cpp, test_core.cpp — SURVEY.md §4), restated as rays through Scene::intersect.
They run against the oracle on CPU and, marked gpu, against the CUDA path through rt_trace_rays."""
import json

import numpy as np
import pytest

MAT = {"diffuse": [0.8, 0.2, 0.2]}
SCREEN = {"dpi": 100, "dimensions": [2, 2], "position": [-1, -1, 0], "observer": [0, 0, 1]}
sph = lambda p, r: {"sphere": {"position": p, "radius": r, "color": MAT}}


def scene(objects):
    return json.dumps({"screen": SCREEN, "objects": objects, "sources": [{"position": [0, 5, 5], "intensity": [10, 10, 10]}]})


CASES = {
    # test_geometry.cpp:17-46: unit sphere, ray (-2,0,0)->+x: t=1, p.x=-1, n=(-1,0,0); (-2,2,0)->+x misses
    "sphere": (scene([sph([0, 0, 0], 1.0)]), [[-2, 0, 0, 1, 0, 0], [-2, 2, 0, 1, 0, 0], [0, 0, 0, 3, 4, 0]]),
    # test_geometry.cpp:57-74: plane x=0 (normal +x) from (-1,0,0): t=1; parallel ray misses
    "plane": (scene([{"halfSpace": {"position": [0, 0, 0], "normal": [2, 0, 0], "color": MAT}}]), [[-1, 0, 0, 1, 0, 0], [-1, 0, 0, 0, 1, 0]]),
    # test_geometry.cpp:89-108: default pokeball, rays from +-y hit different materials (top/bottom), |p|=1
    "pokeball": (scene([{"pokeball": {"position": [0, 0, 0], "radius": 1.0}}]), [[0, 3, 0, 0, -1, 0], [0, -3, 0, 0, 1, 0], [3, 0, 0, -1, 0, 0]]),
    # test_csg.cpp:11-27: union of spheres (-+0.5,0,0) r 0.7, ray (-2,0,0)->+x hits with t<2
    "union": (scene([{"csg": {"operator": "union", "left": sph([-0.5, 0, 0], 0.7), "right": sph([0.5, 0, 0], 0.7)}}]), [[-2, 0, 0, 1, 0, 0]]),
    # test_csg.cpp:31-46: concentric r1 minus r0.5, ray (0,0,-2)->+z hits
    "difference": (scene([{"csg": {"operator": "difference", "left": sph([0, 0, 0], 1.0), "right": sph([0, 0, 0], 0.5)}}]), [[0, 0, -2, 0, 0, 1]]),
}


def check(name, rec):
    hit, t, p, n, front, code = rec[:, 0], rec[:, 1], rec[:, 2:5], rec[:, 5:8], rec[:, 8], rec[:, 9]
    if name == "sphere":
        assert hit[0] == 1 and t[0] == pytest.approx(1.0, abs=1e-5) and p[0, 0] == pytest.approx(-1.0, abs=1e-5)
        assert n[0] == pytest.approx([-1, 0, 0], abs=1e-5) and front[0] == 1
        assert hit[1] == 0 and np.isinf(t[1])
        assert hit[2] == 1 and front[2] == 0 and t[2] == pytest.approx(1.0, abs=1e-5)     # from inside: far root, flipped normal
        assert n[2] == pytest.approx([-0.6, -0.8, 0], abs=1e-5)                          # Ray ctor normalises (3,4,0) (test_core.cpp:130-133)
    elif name == "plane":
        assert hit[0] == 1 and t[0] == pytest.approx(1.0, abs=1e-6) and n[0] == pytest.approx([-1, 0, 0], abs=1e-6)
        assert hit[1] == 0
    elif name == "pokeball":
        assert hit[0] == 1 and hit[1] == 1 and code[0] == 1 and code[1] == 2                # region codes top=1, bottom=2
        assert np.linalg.norm(p[0]) == pytest.approx(1.0, abs=1e-5) and np.linalg.norm(p[1]) == pytest.approx(1.0, abs=1e-5)
        assert code[2] == 5                                                                 # +x is the default button direction
    elif name == "union":
        assert hit[0] == 1 and t[0] < 2.0 and t[0] == pytest.approx(0.8, abs=1e-5)
    elif name == "difference":
        assert hit[0] == 1 and t[0] == pytest.approx(1.0, abs=1e-5) and n[0] == pytest.approx([0, 0, -1], abs=1e-5)


@pytest.mark.parametrize("name", sorted(CASES))
def test_known_answers_oracle(oa, name):
    text, rays = CASES[name]
    check(name, oa.Oracle(text).trace_rays(np.array(rays, float)))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_known_answers_gpu(gpu, name):
    text, rays = CASES[name]
    s = gpu.load_scene_from_json_text(text)
    check(name, gpu.Tracer(s).trace_rays(np.array(rays, float)))


def test_camera_known_answers(oa):
    # test_camera.cpp:18-39: eye (0,0,1), screen P(-1,-1,0) 2x2 dpi 100: centre rays start at the eye, d.z < 0, unit length.
    o = oa.Oracle(scene([sph([0, 0, -3], 1.0)]))
    assert (o.width, o.height) == (200, 200)
    ids = o.render("paper", want_fb=False, want_ids=True)["ids"]
    assert ids[100, 100, 0] == 0 and ids[0, 0, 0] == -1       # centre pixel sees the sphere, the corner does not


def test_recursion_zero_is_black(oa):                         # tracer.cpp:26: depth >= limit at depth 0
    text = json.loads(scene([sph([0, 0, -3], 1.0)])); text["medium"] = {"recursion": 0}; text["background"] = [1, 1, 1]
    fb = oa.Oracle(json.dumps(text)).render("standard", rows=(90, 110))["fb"]
    assert not fb.any()
