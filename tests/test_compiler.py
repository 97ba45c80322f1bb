"""Host scene compiler (csrc/scene_compiler.cpp): flattening, baking of transforms, CSG programs."""
import json
import math

import pytest

MAT = {"diffuse": [1, 0, 0]}
SCREEN = {"dpi": 10, "dimensions": [2, 2], "position": [-1, -1, 0], "observer": [0, 0, 5]}
sph = lambda p, r=1.0: {"sphere": {"position": p, "radius": r, "color": MAT}}
PRIM_SPHERE, PRIM_HALFSPACE, PRIM_POKEBALL, PRIM_PROGRAM = 0, 1, 2, 3
F_XFORMED, F_AFFINE, F_WRAPPED, F_NEVER = 1, 2, 4, 8
OP_LEAF, OP_COMBINE_LEAF, OP_PUSH, OP_COMBINE_POP, OP_XFIX, OP_COMBINE_SPHERES = range(6)


def compile_(rt, objects, **extra):
    return rt.load_scene_from_json_text(json.dumps(dict({"screen": SCREEN, "objects": objects}, **extra)))


def test_objects_keep_json_order(rt):
    s = compile_(rt, [sph([0, 0, 0]), {"halfSpace": {"position": [0, -1, 0], "normal": [0, 2, 0], "color": MAT}},
                      {"pokeball": {"position": [1, 0, 0], "radius": 0.5}}])
    o = s.objects()
    assert [x.kind for x in o] == [PRIM_SPHERE, PRIM_HALFSPACE, PRIM_POKEBALL]
    assert all(x.flags == 0 for x in o)
    assert list(o[1].geom) == pytest.approx([0, 1, 0, -1])        # unit normal, n.p0
    assert s.desc.n_materials == 1 + 1 + 5 and o[2].mat == 2      # pokeball: 5 consecutive region materials


def test_translation_and_rotation_are_baked(rt):
    s = compile_(rt, [{"translation": {"factors": [1, 2, 3], "subject": sph([1, 0, 0], 0.5)}},
                      {"rotation": {"angle": 90, "direction": 2, "subject": sph([1, 0, 0], 0.5)}},
                      {"scaling": {"factors": [2, 2, 2], "subject": sph([1, 0, 0], 0.5)}}])
    o = s.objects()
    assert all(x.kind == PRIM_SPHERE and x.flags == (F_XFORMED | F_WRAPPED) for x in o)
    assert list(o[0].geom) == pytest.approx([2, 2, 3, 0.5])
    assert list(o[1].geom) == pytest.approx([0, 1, 0, 0.5], abs=1e-6)   # Rz(90) maps +x to +y
    assert list(o[2].geom) == pytest.approx([2, 0, 0, 1.0])            # uniform scale: still a sphere
    assert s.desc.n_affines == 0


def test_non_uniform_scaling_keeps_an_affine_map(rt):
    s = compile_(rt, [{"scaling": {"factors": [2, 1, 1], "subject": sph([1, 0, 0], 0.5)}}])
    o = s.objects()[0]
    assert o.flags == (F_XFORMED | F_WRAPPED | F_AFFINE) and o.xf == 0 and s.desc.n_affines == 1
    m = s.desc.affines[0].m
    assert [m[0][0], m[1][1], m[2][2]] == pytest.approx([0.5, 1, 1])
    assert list(o.geom) == pytest.approx([1, 0, 0, 0.5])              # geometry stays in the local frame


def test_degenerate_scaling_never_hits(rt):                            # transform.cpp:103
    s = compile_(rt, [{"scaling": {"factors": [1e-7, 1, 1], "subject": sph([0, 0, 0])}}])
    assert s.objects()[0].flags & F_NEVER


def test_halfspace_under_rotation(rt):
    s = compile_(rt, [{"rotation": {"angle": 90, "direction": 0, "subject": {"halfSpace": {"position": [0, 1, 0], "normal": [0, 1, 0], "color": MAT}}}}])
    g = list(s.objects()[0].geom)                                      # Rx(90): +y -> +z, plane y>=1 -> z>=1
    assert g == pytest.approx([0, 0, 1, 1], abs=1e-6)


def test_pokeball_axes_follow_rotation(rt):
    s = compile_(rt, [{"rotation": {"angle": 90, "direction": 2, "subject": {"pokeball": {"position": [0, 0, 0], "radius": 1, "button_dir": [0, 0, 2]}}}}])
    pb = s.desc.pokeballs[0]
    assert list(pb.up) == pytest.approx([-1, 0, 0], abs=1e-6)          # Rz(90): +y -> -x
    assert list(pb.btn_dir) == pytest.approx([0, 0, 1], abs=1e-6)
    assert pb.cos_outer == pytest.approx(math.cos(0.28)) and pb.cos_inner == pytest.approx(math.cos(0.22))
    assert pb.belt_half == pytest.approx(0.06)


def decode(w):
    return (w & 15, (w >> 4) & 3, w >> 8)


def test_left_spine_union_needs_no_stack_and_fuses(rt):
    s = compile_(rt, [{"union": [sph([i, 0, 0]) for i in range(5)]}])
    o = s.objects()[0]
    assert o.kind == PRIM_PROGRAM and o.flags == 0
    ops = [decode(w) for w in s.ops()]
    assert ops[0] == (OP_LEAF, 0, 0)
    assert ops[1][0] == OP_COMBINE_SPHERES and ops[1][2] == (1 | (4 << 14)) and len(ops) == 2
    assert s.desc.n_leaves == 5


def test_right_nested_csg_uses_push_pop(rt):
    inner = {"csg": {"operator": "difference", "left": sph([0, 0, 0]), "right": sph([0.5, 0, 0], 0.5)}}
    s = compile_(rt, [{"csg": {"operator": "intersection", "left": sph([0, 0, 0]), "right": inner}}])
    ops = [decode(w)[:2] for w in s.ops()]
    assert ops == [(OP_LEAF, 0), (OP_PUSH, 0), (OP_LEAF, 0), (OP_COMBINE_LEAF, 2), (OP_COMBINE_POP, 1)]


def test_transform_inside_csg_emits_xfix_only_over_csg(rt):
    inner = {"union": [sph([0, 0, 0]), sph([1, 0, 0])]}
    s = compile_(rt, [{"union": [{"rotation": {"angle": 10, "direction": 1, "subject": inner}},
                                 {"translation": {"factors": [0, 1, 0], "subject": sph([0, 0, 0])}}]}])
    ops = [decode(w) for w in s.ops()]
    assert [o[0] for o in ops] == [OP_LEAF, OP_COMBINE_LEAF, OP_XFIX, OP_COMBINE_LEAF]
    assert ops[2][2] == 0 and s.desc.n_fallbacks == 1                  # rotation over CSG: zero normals -> fallback 0
    assert all(l.flags & F_XFORMED for l in s.leaves())


def test_wrapped_program_flags(rt):
    s = compile_(rt, [{"translation": {"factors": [0, -0.3, 0], "subject": {"union": [sph([0, 0, 0]), sph([1, 0, 0])]}}}])
    o = s.objects()[0]
    assert o.kind == PRIM_PROGRAM and o.flags == (F_WRAPPED | F_XFORMED)
    assert [decode(w)[0] for w in s.ops()] == [OP_LEAF, OP_COMBINE_LEAF]   # the stripped top transform emits no XFIX
    assert list(s.leaves()[0].geom) == pytest.approx([0, -0.3, 0, 1])


def test_limits(rt):
    deep = sph([0, 0, 0])
    for _ in range(12):                                                    # right-nesting deeper than RT_MAX_CSG_STACK
        deep = {"csg": {"operator": "union", "left": sph([0, 0, 0]), "right": {"csg": {"operator": "union", "left": deep, "right": sph([1, 0, 0])}}}}
    with pytest.raises(rt.RtError) as e:
        compile_(rt, [deep])
    assert e.value.code == -4
    with pytest.raises(rt.RtError) as e:
        compile_(rt, [sph([0, 0, 0])], medium={"recursion": 99})
    assert e.value.code == -4


def test_example_scenes_compile(rt):
    from conftest import scene_path
    s = rt.load_scene_from_json(scene_path("snorlax_4k"))
    assert (s.width, s.height) == (3840, 2160) and s.desc.n_objects == 15 and s.desc.n_leaves == 20 and s.desc.n_lights == 5
    a = rt.load_scene_from_json(scene_path("synthetic_a_8k"))
    assert (a.width, a.height) == (7680, 4320) and a.desc.n_objects == 1 + 1000 + 64 + 32 and a.desc.recursion == 8
    b = rt.load_scene_from_json(scene_path("synthetic_b_16k"))
    assert (b.width, b.height) == (16384, 16384)
