"""Loader parity: the assertions of the reference's tests/test_json_loader.cpp, restated
against BOTH the product loader (csrc/json_loader.cpp through the C ABI) and the oracle's loader."""
import json

import pytest

MAT = {"diffuse": [1, 0, 0]}
BASE = {"screen": {"dpi": 100, "dimensions": [2, 2], "position": [-1, -1, 0], "observer": [0, 0, 5]},
        "objects": [{"sphere": {"position": [0, 0, 0], "radius": 1.0, "color": MAT}}]}


def loaders(rt, oa):
    def product(text):
        s = rt.load_scene_from_json_text(text)
        return {"n_objects": int(s.desc.n_objects), "W": s.width, "H": s.height, "n_lights": int(s.desc.n_lights),
                "recursion": int(s.desc.recursion)}

    def oracle(text):
        o = oa.Oracle(text)
        return {"n_objects": o.n_objects, "W": o.width, "H": o.height, "n_lights": o.n_lights, "recursion": o.recursion}
    return [("product", product, rt.RtError), ("oracle", oracle, oa.OracleError)]


def test_valid_scene_loads(rt, oa):                      # test_json_loader.cpp:42-49
    for name, load, _ in loaders(rt, oa):
        r = load(json.dumps(BASE))
        assert r["n_objects"] == 1 and r["W"] == 200 and r["H"] == 200, name


def test_default_dpi_is_72(rt, oa):                      # test_json_loader.cpp:52-66
    s = json.loads(json.dumps(BASE)); del s["screen"]["dpi"]
    for name, load, _ in loaders(rt, oa):
        r = load(json.dumps(s))
        assert (r["W"], r["H"]) == (144, 144), name


def test_defaults_without_medium(rt, oa):                # scene.h:43-47
    for name, load, _ in loaders(rt, oa):
        assert load(json.dumps(BASE))["recursion"] == 5, name


def test_invalid_json_throws(rt, oa):                    # test_json_loader.cpp:69-74
    for name, load, Err in loaders(rt, oa):
        with pytest.raises(Err) as e:
            load("{ invalid json }")
        assert "JSON parse error" in str(e.value), name


@pytest.mark.parametrize("mutate,needle", [
    (lambda s: s["objects"][0]["sphere"].__setitem__("color", [1, 0, 0]), "color must be an object"),          # :77-98
    (lambda s: s["objects"][0]["sphere"].pop("radius"), "sphere requires"),
    (lambda s: s["objects"].__setitem__(0, {"sphere": {}, "halfSpace": {}}), "one-entry object"),
    (lambda s: s["objects"].__setitem__(0, {"cube": {}}), "unknown object kind: cube"),
    (lambda s: s["screen"].pop("position"), "screen.position is required"),
    (lambda s: s["screen"].pop("observer"), "screen.observer is required"),
    (lambda s: s["screen"].__setitem__("dimensions", [1, 2, 3]), "screen.dimensions must be [Lx, Ly]"),
    (lambda s: s.__setitem__("sources", {"position": [0, 0, 0]}), "'sources' must be an array"),
    (lambda s: s.__setitem__("sources", [{"position": [0, 0, 0]}]), "each source needs"),
    (lambda s: s.__setitem__("objects", {}), "'objects' must be an array"),
    (lambda s: s.__setitem__("background", [1, 2]), "Expected color array[3]"),
    (lambda s: s["objects"].__setitem__(0, {"rotation": {"angle": 30, "direction": 5, "subject": s["objects"][0]}}),
     "rotation direction must be 0 (X), 1 (Y), or 2 (Z)"),                                                   # :177-210
    (lambda s: s["objects"].__setitem__(0, {"csg": {"operator": "xor", "left": s["objects"][0], "right": s["objects"][0]}}),
     "csg.operator must be union/intersection/difference"),
    (lambda s: s["objects"].__setitem__(0, {"difference": [s["objects"][0]]}), "difference array must have at least 2"),
    (lambda s: s["objects"].__setitem__(0, {"union": []}), "CSG array must be a non-empty array"),
    (lambda s: s["objects"].__setitem__(0, {"translation": {"subject": s["objects"][0]}}), "translation requires 'factors' and 'subject'"),
    (lambda s: s["objects"].__setitem__(0, {"pokeball": {"radius": 1}}), "pokeball requires 'position' and 'radius'."),
])
def test_processing_errors(rt, oa, mutate, needle):
    s = json.loads(json.dumps(BASE)); mutate(s)
    for name, load, Err in loaders(rt, oa):
        with pytest.raises(Err) as e:
            load(json.dumps(s))
        assert "JSON processing error" in str(e.value) and needle in str(e.value), (name, str(e.value))


def test_every_node_kind_loads(rt, oa):                  # test_json_loader.cpp:130-133,171-174 and json_loader.cpp:409-434
    sph = BASE["objects"][0]
    s = json.loads(json.dumps(BASE))
    s["objects"] = [
        {"halfSpace": {"position": [0, -1, 0], "normal": [0, 1, 0], "color": MAT}},
        {"pokeball": {"position": [0, 0, 0], "radius": 1}},
        {"rotation": {"angle": 45, "direction": 2, "subject": sph}},
        {"scaling": {"factors": [1, 2, 3], "subject": sph}},
        {"translation": {"factors": [1, 2, 3], "subject": sph}},
        {"csg": {"operator": "difference", "left": sph, "right": sph}},
        {"union": [sph, sph, sph]}, {"intersection": [sph]}, {"difference": [sph, sph, sph]},
    ]
    s["sources"] = [{"position": [1, 2, 3], "intensity": [1, 1, 1]}]
    s["medium"] = {"ambient": [0.1, 0.1, 0.1], "index": 1.2, "recursion": 3}
    for name, load, _ in loaders(rt, oa):
        r = load(json.dumps(s))
        assert r["n_objects"] == 9 and r["n_lights"] == 1 and r["recursion"] == 3, name


def test_missing_file(rt):
    with pytest.raises(rt.RtError) as e:
        rt.load_scene_from_json("/nonexistent/scene.json")
    assert "Cannot open JSON file" in str(e.value)


def test_material_rules(rt):                             # json_loader.cpp:83-123: means of RGB triples, defaults
    s = json.loads(json.dumps(BASE))
    s["medium"] = {"ambient": [0.5, 0.5, 0.5]}
    s["objects"][0]["sphere"]["color"] = {"diffuse": [0.1, 0.2, 0.3], "ambient": [0.2, 0.4, 0.6], "specular": [0.3, 0.6, 0.9],
                                          "reflected": [0.1, 0.2, 0.3], "refracted": [0.9, 0.6, 0.3], "shininess": 7, "kd": 0.1, "ks": 0.9}
    s["objects"][0]["sphere"]["index"] = 1.5
    m = rt.load_scene_from_json_text(json.dumps(s)).materials()[0]
    assert list(m.albedo) == pytest.approx([0.1, 0.2, 0.3])
    assert list(m.ambient) == pytest.approx([0.1, 0.2, 0.3])          # pre-multiplied by medium.ambient (shading.cpp:39)
    assert (m.ks, m.kr, m.kt) == pytest.approx((0.6, 0.2, 0.6))
    assert (m.shininess, m.ior, m.kd) == pytest.approx((7, 1.5, 1.0))   # kd/ks keys in JSON are ignored
    d = rt.load_scene_from_json_text(json.dumps(BASE)).materials()[0]
    assert (d.ks, d.kr, d.kt, d.shininess, d.ior) == (0, 0, 0, 1, 1)


def test_number_syntax_is_parsed_like_a_correctly_rounding_reader(rt):
    """scenes/number_syntax.json stresses the number grammar (exponents, -0, 30-digit mantissas, duplicate keys, escapes).
    The reference reads doubles with nlohmann/json (strtod: correctly rounded), as does Python's json; the product's reader
    must produce the same doubles, seen here through the float32 geometry and the double camera of the compiled scene."""
    import os
    import numpy as np
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes", "number_syntax.json")
    want = json.load(open(path))                          # duplicate keys: the last one wins, as in nlohmann::json
    s = rt.load_scene_from_json(path)
    assert (s.width, s.height) == (96, 72)                # round(2.0E+0 * 4.8e1), round(15e-1 * 4.8e1)  (camera.h:19-21)
    o = s.objects()
    sph = want["objects"][1]["sphere"]
    assert sph["radius"] == 0.65
    assert list(o[1].geom) == [float(np.float32(v)) for v in sph["position"] + [sph["radius"]]]
    assert list(o[2].geom)[3] == float(np.float32(4.9999999999999994e-1))
    assert list(o[0].geom)[:3] == [0.0, 1.0, 0.0]         # 123456789012345678901234567890e-29 normalised
    lights = [s.desc.lights[i] for i in range(s.desc.n_lights)]
    assert [float(v) for v in lights[1].intensity] == [float(np.float32(v)) for v in want["sources"][1]["intensity"]]
