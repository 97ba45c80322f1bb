#!/usr/bin/env python3
"""Generate every scene file the tests and the bench use, into scenes/ (committed).

* The three example scenes of the reference (examples/*.json) are re-serialised from
  /root/reference when it is present (this container), together with the derived variants
  SURVEY.md §8(d) names: only the `screen` block (and, for snorlax, `medium.recursion` and two
  extra lights) changes.  Resolution can only be set through the JSON (nx = round(Lx*dpi),
  camera.h:19-21), so every non-native config is a derived file.
* Synthetic-A / Synthetic-B are deterministic (seeded numpy) and use only reference-format nodes.

The GPU box has no /root/reference: the generated files are what travels.
"""
from __future__ import annotations

import copy
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "scenes")
REF_EXAMPLES = "/root/reference/examples"

SCREEN_4K = {"dpi": 720, "dimensions": [5.333333333333333, 3], "position": [-2.6666666666666665, -1.5, 0]}


def dump(name: str, scene: dict) -> None:
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, name), "w") as f:
        json.dump(scene, f, separators=(",", ":"))
        f.write("\n")


def with_screen(scene: dict, **kw) -> dict:
    s = copy.deepcopy(scene)
    s["screen"].update(kw)
    return s


def scaled(scene: dict, dpi: int) -> dict:
    return with_screen(scene, dpi=dpi)


def examples() -> None:
    if not os.path.isdir(REF_EXAMPLES):
        print("reference examples absent: keeping the committed scenes/", file=sys.stderr)
        return
    load = lambda n: json.load(open(os.path.join(REF_EXAMPLES, n)))
    pk, pg, sn = load("pokeballs.json"), load("penguin.json"), load("snorlax.json")
    dump("pokeballs.json", pk)                       # config 1: native 1024x768
    dump("pokeballs_small.json", scaled(pk, 64))     # 256x192, fast parity case
    dump("penguin.json", pg)                         # config 2: native 1200x900 (standard and --paper)
    dump("penguin_small.json", scaled(pg, 80))       # 320x240
    dump("penguin_4k.json", with_screen(pg, **SCREEN_4K))
    dump("snorlax.json", sn)
    dump("snorlax_small.json", scaled(sn, 60))       # 240x180
    # config 3: 3840x2160, recursion 5, 5 lights (3 originals + 2 fixed extras)
    s4 = with_screen(sn, **SCREEN_4K)
    s4["medium"]["recursion"] = 5
    s4["sources"] = s4["sources"] + [{"position": [0, 5, 3], "intensity": [10, 10, 10]},
                                     {"position": [0, 2, 9], "intensity": [15, 15, 15]}]
    dump("snorlax_4k.json", s4)
    small = copy.deepcopy(s4); small["screen"].update({"dpi": 60})   # 320x180, same scene
    dump("snorlax_4k_small.json", small)
    # variant that exercises the bounce path: reflective belly sphere inside the CSG union
    refl = copy.deepcopy(s4)
    belly = refl["objects"][-1]["translation"]["subject"]["union"][8]["sphere"]
    belly["color"]["reflected"] = [0.3, 0.3, 0.3]
    dump("snorlax_4k_reflect.json", refl)
    rs = copy.deepcopy(refl); rs["screen"].update({"dpi": 60})
    dump("snorlax_4k_reflect_small.json", rs)


def color(rng, reflect=0.0, refract=0.0, spec=0.3, shin=32):
    c = {"diffuse": [round(float(v), 3) for v in rng.uniform(0.15, 0.95, 3)],
         "ambient": [0.1, 0.1, 0.1], "specular": [spec] * 3, "shininess": shin}
    if reflect:
        c["reflected"] = [reflect] * 3
    if refract:
        c["refracted"] = [refract] * 3
    return c


def synthetic_a(dpi: int, n_spheres=1000, n_pokeballs=64, n_csg=32, recursion=8) -> dict:
    """Config 4: 1,000 spheres + 64 pokeballs + half-space + 32 nested CSG (difference of intersection) under transforms."""
    rng = np.random.default_rng(20261104)
    objs = [{"halfSpace": {"position": [0, -1, 0], "normal": [0, 1, 0],
                           "color": {"diffuse": [0.55, 0.55, 0.5], "ambient": [0.1, 0.1, 0.1], "reflected": [0.15] * 3}}}]
    for i in range(n_spheres):
        c = [float(rng.uniform(-8, 8)), float(rng.uniform(-0.5, 6)), float(rng.uniform(-12, -2))]
        r = float(rng.uniform(0.08, 0.35))
        u = rng.uniform()
        col = color(rng, reflect=0.4) if u < 0.30 else color(rng, refract=0.6) if u < 0.45 else color(rng)
        node = {"position": [round(v, 4) for v in c], "radius": round(r, 4), "color": col}
        if u >= 0.30 and u < 0.45:
            node["index"] = 1.5
        objs.append({"sphere": node})
    for i in range(n_pokeballs):
        gx, gz = i % 8, i // 8
        c = [-7 + 2 * gx + float(rng.uniform(-0.3, 0.3)), -0.7, -3.0 - 1.1 * gz + float(rng.uniform(-0.2, 0.2))]
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        objs.append({"pokeball": {"position": [round(v, 4) for v in c], "radius": 0.3,
                                  "button_dir": [round(float(v), 4) for v in d]}})
    for i in range(n_csg):
        c = np.array([float(rng.uniform(-6, 6)), float(rng.uniform(0.5, 4)), float(rng.uniform(-9, -3))])
        a = {"sphere": {"position": [0, 0, 0], "radius": 0.6, "color": color(rng, reflect=0.2 if i % 3 == 0 else 0)}}
        b = {"sphere": {"position": [0.45, 0, 0], "radius": 0.6, "color": color(rng)}}
        hole = {"sphere": {"position": [0.22, 0, 0.45], "radius": 0.3, "color": color(rng)}}
        node = {"difference": [{"intersection": [a, b]}, hole]}
        kind = i % 4
        if kind == 1:
            node = {"rotation": {"angle": float(round(rng.uniform(0, 360), 2)), "direction": int(i % 3), "subject": node}}
        elif kind == 2:
            node = {"scaling": {"factors": [1.5, 1.5, 1.5], "subject": node}}
            c = c / 1.0
        elif kind == 3:
            node = {"scaling": {"factors": [1.0, 1.6, 0.8], "subject": node}}
        node = {"translation": {"factors": [round(float(v), 4) for v in c], "subject": node}}
        objs.append(node)
    lights = [{"position": [-6, 9, 4], "intensity": [40, 40, 40]}, {"position": [6, 8, 5], "intensity": [35, 35, 30]},
              {"position": [0, 10, -6], "intensity": [30, 30, 35]}, {"position": [-8, 5, -10], "intensity": [20, 20, 25]},
              {"position": [8, 4, -1], "intensity": [25, 20, 20]}]
    return {"screen": {"dpi": dpi, "dimensions": [5.333333333333333, 3], "position": [-2.6666666666666665, 0.5, 3], "observer": [0, 2, 8]},
            "medium": {"ambient": [0.3, 0.3, 0.35], "index": 1.0, "recursion": recursion},
            "background": [0.35, 0.55, 0.85], "sources": lights, "objects": objs}


def synthetic_b(dpi: int, recursion=6) -> dict:
    """Config 5: refraction-heavy glass lattice (medium index 1.0, objects index 1.5); kr and kt both > 0 -> binary ray trees."""
    rng = np.random.default_rng(20261105)
    objs = [{"halfSpace": {"position": [0, -1.6, 0], "normal": [0, 1, 0],
                           "color": {"diffuse": [0.6, 0.6, 0.55], "ambient": [0.1, 0.1, 0.1]}}}]
    for x, y, z, col in [(-2.2, 0.3, -6.5, [0.9, 0.2, 0.2]), (2.2, 0.3, -6.5, [0.2, 0.9, 0.2]),
                         (-0.8, 1.8, -7.0, [0.2, 0.3, 0.9]), (0.9, -0.8, -6.0, [0.9, 0.8, 0.2])]:
        objs.append({"sphere": {"position": [x, y, z], "radius": 0.8,
                                "color": {"diffuse": col, "ambient": [0.1, 0.1, 0.1], "specular": [0.2] * 3, "shininess": 16}}})
    glass = {"diffuse": [0.05, 0.05, 0.06], "ambient": [0.02, 0.02, 0.02], "specular": [0.6] * 3, "shininess": 64,
             "refracted": [0.9, 0.9, 0.9], "reflected": [0.1, 0.1, 0.1]}
    for ix in range(5):
        for iy in range(5):
            for iz in range(3):
                c = [-1.6 + 0.8 * ix, -1.2 + 0.8 * iy, -1.0 - 0.9 * iz]
                objs.append({"sphere": {"position": [round(v, 3) for v in c], "radius": 0.33, "color": glass, "index": 1.5}})
    lights = [{"position": [4, 6, 6], "intensity": [40, 40, 40]}, {"position": [-5, 5, 4], "intensity": [30, 30, 35]},
              {"position": [0, 8, -3], "intensity": [25, 25, 25]}]
    return {"screen": {"dpi": dpi, "dimensions": [4, 4], "position": [-2, -2, 0], "observer": [0, 0, 5]},
            "medium": {"ambient": [0.25, 0.25, 0.3], "index": 1.0, "recursion": recursion},
            "background": [0.1, 0.12, 0.2], "sources": lights, "objects": objs}


def quirks() -> dict:
    """A small scene built to hit every node kind and the CSG / transform quirks of SURVEY.md §3.4."""
    m = lambda d, **kw: dict({"diffuse": d, "ambient": [0.2, 0.2, 0.2], "specular": [0.3] * 3, "shininess": 20}, **kw)
    sph = lambda p, r, col, **kw: {"sphere": dict({"position": p, "radius": r, "color": col}, **kw)}
    objs = [
        {"halfSpace": {"position": [0, -1.5, 0], "normal": [0, 2, 0], "color": m([0.5, 0.6, 0.5], reflected=[0.2] * 3)}},
        sph([-2.4, 0.9, -2], 0.55, m([0.8, 0.2, 0.2], reflected=[0.5] * 3)),
        sph([2.4, 0.9, -2], 0.55, m([0.1, 0.1, 0.12], refracted=[0.8] * 3, reflected=[0.1] * 3), index=1.5),
        {"pokeball": {"position": [0, 1.0, -1.5], "radius": 0.5, "button_dir": [0.3, 0.2, 1.0]}},
        {"rotation": {"angle": 35, "direction": 2, "subject": {"pokeball": {"position": [1.3, -0.2, -1.0], "radius": 0.45,
                                                                           "button_dir": [0, 0, 1], "belt_half": 0.1}}}},
        {"csg": {"operator": "difference", "left": sph([-1.2, -0.4, -1], 0.6, m([0.9, 0.6, 0.1])),
                 "right": sph([-1.0, -0.2, -0.6], 0.4, m([0.1, 0.6, 0.9]))}},
        {"translation": {"factors": [0.1, -0.6, 0.3], "subject": {"intersection": [
            sph([0, 0, -1], 0.6, m([0.2, 0.8, 0.3])), sph([0.35, 0.1, -0.9], 0.5, m([0.8, 0.3, 0.8])),
            {"halfSpace": {"position": [0, 0.25, 0], "normal": [0, -1, 0], "color": m([0.9, 0.9, 0.2])}}]}}},
        {"scaling": {"factors": [1.6, 0.6, 1.0], "subject": sph([1.2, 2.6, -2.5], 0.5, m([0.3, 0.3, 0.9], reflected=[0.3] * 3))}},
        {"translation": {"factors": [-2.2, -0.9, 0.2], "subject": {"scaling": {"factors": [0.7, 0.7, 0.7], "subject": {
            "union": [sph([0, 0, 0], 0.5, m([0.9, 0.4, 0.4])),
                      {"rotation": {"angle": 50, "direction": 1, "subject": {"csg": {"operator": "difference",
                          "left": sph([0.6, 0.2, 0], 0.45, m([0.4, 0.9, 0.4])),
                          "right": {"translation": {"factors": [0.1, 0, 0.3], "subject": sph([0.6, 0.2, 0], 0.3, m([0.4, 0.4, 0.9]))}}}}}},
                      {"scaling": {"factors": [1, 2, 1], "subject": sph([-0.5, 0.2, 0.1], 0.25, m([0.9, 0.9, 0.3]))}}]}}}}},
        {"union": [sph([2.3, -0.8, -0.5], 0.4, m([0.7, 0.7, 0.7])),
                   {"translation": {"factors": [0, 0, 0], "subject": {"halfSpace": {"position": [0, -1.0, 0], "normal": [0, -1, 0], "color": m([0.3, 0.3, 0.3])}}}}]},
    ]
    return {"screen": {"dpi": 80, "dimensions": [4, 3], "position": [-2, -1.5, 1], "observer": [0, 0.3, 6]},
            "medium": {"ambient": [0.4, 0.4, 0.45], "index": 1.0, "recursion": 4},
            "background": [0.3, 0.45, 0.7],
            "sources": [{"position": [4, 6, 6], "intensity": [30, 30, 30]}, {"position": [-5, 4, 3], "intensity": [20, 20, 25]},
                        {"position": [0, 0.2, 2.5], "intensity": [6, 6, 6]}],
            "objects": objs}


def quirks_far() -> dict:
    """quirks seen from 36 units away (primary t ~ 35-40: the FP64 epsilons 1e-6 of the CSG comparator and of the
    secondary-ray offsets are below one FP32 ulp there), plus CSG nodes whose children have COINCIDENT faces:
    identical spheres under difference / intersection / union, a zero-thickness slab between two half-spaces on one
    plane, two spheres touching in a point, and a reflective + a refractive sphere far from the eye."""
    sc = quirks()
    m = lambda d, **kw: dict({"diffuse": d, "ambient": [0.2, 0.2, 0.2], "specular": [0.3] * 3, "shininess": 20}, **kw)
    sph = lambda p, r, col, **kw: {"sphere": dict({"position": p, "radius": r, "color": col}, **kw)}
    hs = lambda p, n, col: {"halfSpace": {"position": p, "normal": n, "color": col}}
    sc["objects"] += [
        {"csg": {"operator": "difference", "left": sph([-0.2, 2.1, -2.0], 0.4, m([0.9, 0.2, 0.6])), "right": sph([-0.2, 2.1, -2.0], 0.4, m([0.2, 0.9, 0.6]))}},
        {"intersection": [sph([-1.3, 2.0, -2.0], 0.4, m([0.9, 0.5, 0.2])), sph([-1.3, 2.0, -2.0], 0.4, m([0.2, 0.5, 0.9]))]},
        {"union": [sph([0.35, -1.0, 0.2], 0.3, m([0.9, 0.9, 0.9])), sph([0.35, -1.0, 0.2], 0.3, m([0.1, 0.1, 0.1])),
                   sph([0.95, -1.0, 0.2], 0.3, m([0.6, 0.2, 0.2]))]},                      # the third one touches the pair in one point
        {"intersection": [sph([-0.6, -1.0, 0.4], 0.35, m([0.3, 0.7, 0.7])), hs([0, -1.0, 0], [0, 1, 0], m([0.7, 0.7, 0.3])),
                          hs([0, -1.0, 0], [0, -1, 0], m([0.7, 0.3, 0.7]))]},               # slab of zero thickness
        {"csg": {"operator": "difference", "left": sph([1.5, 1.9, -1.5], 0.45, m([0.8, 0.8, 0.2])),
                 "right": hs([0, 1.9, 0], [0, 1, 0], m([0.2, 0.2, 0.8]))}},                  # cut through the centre: coplanar with nothing, control
        sph([-2.6, -0.6, -3.0], 0.5, m([0.7, 0.7, 0.75], reflected=[0.6] * 3)),
        sph([1.9, 0.55, -3.2], 0.5, m([0.1, 0.1, 0.1], refracted=[0.85] * 3, reflected=[0.1] * 3), index=1.5),
    ]
    sc["screen"] = {"dpi": 80, "dimensions": [4, 3], "position": [-2, -1.2, 1], "observer": [0, 0.3, 37]}
    return sc


def big_floor() -> dict:
    """One huge sphere (the common r = 1000 'floor') plus many small ones: the candidate scan's rounding margin must stay
    per primitive (a global max(|c|^2 + r^2) would make every small sphere a candidate of every ray)."""
    rng = np.random.default_rng(20261105)
    m = lambda d, **kw: dict({"diffuse": d, "ambient": [0.15] * 3, "specular": [0.3] * 3, "shininess": 16}, **kw)
    objs = [{"sphere": {"position": [0, -1001.0, -4], "radius": 1000.0, "color": m([0.5, 0.55, 0.5])}}]
    for _ in range(40):
        p = [float(rng.uniform(-3, 3)), float(rng.uniform(-0.8, 1.6)), float(rng.uniform(-7, -1.5))]
        objs.append({"sphere": {"position": p, "radius": float(rng.uniform(0.12, 0.4)), "color": m([float(v) for v in rng.uniform(0.2, 0.9, 3)])}})
    return {"screen": {"dpi": 60, "dimensions": [4, 3], "position": [-2, -1.5, 0], "observer": [0, 0.2, 5]},
            "medium": {"ambient": [0.3, 0.3, 0.3], "index": 1.0, "recursion": 2}, "background": [0.2, 0.3, 0.5],
            "sources": [{"position": [4, 6, 5], "intensity": [40, 40, 40]}, {"position": [-4, 3, 2], "intensity": [15, 15, 18]}],
            "objects": objs}


def main() -> None:
    examples()
    dump("quirks_far.json", quirks_far())
    dump("big_floor.json", big_floor())
    dump("synthetic_a_8k.json", synthetic_a(1440))          # 7680x4320, recursion 8
    dump("synthetic_a_small.json", synthetic_a(60))          # 320x180, same scene: full-frame CPU parity
    dump("synthetic_a_2k.json", synthetic_a(360))           # 1920x1080, same scene: profiling size
    dump("synthetic_a_tiny.json", synthetic_a(45, n_spheres=120, n_pokeballs=16, n_csg=8, recursion=5))  # CPU-test sized
    dump("synthetic_b_16k.json", synthetic_b(4096))         # 16384x16384, recursion 6
    dump("synthetic_b_4k.json", synthetic_b(1024))          # 4096x4096 single-GPU sized
    dump("synthetic_b_1k.json", synthetic_b(256))          # 1024x1024, same scene: profiling size
    dump("synthetic_b_small.json", synthetic_b(48))         # 192x192
    dump("quirks.json", quirks())
    print("scenes written to", OUT)


if __name__ == "__main__":
    main()
