"""The device's candidate scan (csrc/rt_device.cuh sphere_mask) is conservative: it uses the cheap expanded quadratic
of Sphere::intersect (geometry.cpp:15-22) with a rounding margin, and every candidate is re-tested with the exact
cancellation-free discriminant (sphere_disc).  Host-side checks (no GPU):
  * the mask table the library builds is (c, |c|^2 - r^2 - margin (|c|^2 + r^2)) of the geometry arrays: the rounding
    margin is per primitive (one r = 1000 "floor" sphere must not loosen the mask of the small ones);
  * an FP32 emulation of both forms on adversarial (near-tangent, far-from-origin, badly normalised) rays never rejects
    a primitive the exact form accepts, and the observed error uses a small part of the margin.
FMA is emulated as round32(a*b + c) in float64 (the product of two float32 is exact in float64)."""
import numpy as np
import pytest

from conftest import scene_path

F = np.float32
MARGIN = 1.0 / 65536.0          # kMaskMargin


def fma(a, b, c):
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F)


def dot3(a, b):
    return fma(a[0], b[0], fma(a[1], b[1], (a[2] * b[2]).astype(F)))


def exact_disc(c, r, o, d):
    """sphere_disc: r^2 - |oc - (oc.d) d|^2"""
    oc = [(o[k] - c[k]).astype(F) for k in range(3)]
    hb = dot3(oc, d)
    l = [fma(-hb, d[k], oc[k]) for k in range(3)]
    return fma(-l[0], l[0], fma(-l[1], l[1], fma(-l[2], l[2], (r * r).astype(F))))


def cheap_value_and_bound(c, w, o, d):
    """mask_miss: hb^2 - s against |o|^2 (1 - margin); the primitive's share of the margin is inside w"""
    oo, od = dot3(o, o), dot3(o, d)
    bound = (fma(F(-MARGIN) * np.ones_like(oo), oo, oo) - F(1e-30)).astype(F)
    m2o = [(F(-2.0) * o[k]).astype(F) for k in range(3)]
    hb = fma(-c[0], d[0], fma(-c[1], d[1], fma(-c[2], d[2], od)))
    s = fma(c[0], m2o[0], fma(c[1], m2o[1], fma(c[2], m2o[2], w)))
    return fma(hb, hb, -s), bound, oo


@pytest.mark.parametrize("name", ["snorlax_small", "synthetic_a_small", "synthetic_b_small", "pokeballs_small", "quirks"])
def test_mask_table_mirrors_the_geometry(rt, name):
    sc = rt.load_scene_from_json(scene_path(name))
    masks, scale = sc.mask_table()
    order, runs, clusters = sc.device_layout()
    objs, leaves = sc.objects(), sc.leaves()
    assert len(masks) == len(objs) + len(leaves) + len(clusters)
    geom = [objs[i].geom[:] for i in order.tolist()] + [p.geom[:] for p in leaves] + clusters.tolist()
    best = 0.0
    for m, g in zip(masks.tolist(), geom):
        g = np.array(g, np.float32).astype(np.float64)
        assert m[:3] == g[:3].astype(np.float32).tolist()
        cc, rr = np.dot(g[:3], g[:3]), g[3] * g[3]
        want = float(np.float32((cc - rr) - MARGIN * (cc + rr)))
        assert m[3] == want or (np.isnan(m[3]) and np.isnan(want))
        if np.isfinite(np.dot(g[:3], g[:3]) + g[3] * g[3]):
            best = max(best, float(np.dot(g[:3], g[:3]) + g[3] * g[3]))
    assert scale >= 0.0 and scale <= best * 1.00001 + 1e-30      # (half-spaces / affine leaves are left out of the scale)


@pytest.mark.parametrize("extent, rmax, seed", [(1.0, 0.5, 1), (20.0, 9.0, 2), (20.0, 0.05, 3), (300.0, 2.0, 4), (5000.0, 0.3, 5)])
def test_cheap_form_never_rejects_what_the_exact_form_accepts(extent, rmax, seed):
    rng = np.random.default_rng(seed)
    n = 400_000
    c = [rng.uniform(-extent, extent, n).astype(F) for _ in range(3)]
    r = rng.uniform(rmax * 0.01, rmax, n).astype(F)
    o = [rng.uniform(-extent, extent, n).astype(F) for _ in range(3)]
    # lines at distance r * (1 + tiny) from the centre: near-tangent, both sides of the silhouette
    oc = np.array([c[k].astype(np.float64) - o[k].astype(np.float64) for k in range(3)])
    L = np.linalg.norm(oc, axis=0); e = oc / L
    v = np.cross(e.T, rng.normal(size=(n, 3))).T; v /= np.linalg.norm(v, axis=0)
    delta = rng.choice([-1e-2, -1e-4, -1e-6, -1e-7, 0.0, 1e-7, 1e-6, 1e-4, 1e-2], n) * rng.uniform(0, 1, n)
    sin_phi = np.clip(r.astype(np.float64) * (1.0 + delta) / L, 0.0, 1.0)       # origins inside the sphere: any line hits
    dd = e * np.sqrt(1.0 - sin_phi ** 2) + v * sin_phi
    dd *= 1.0 + rng.uniform(-1.0, 1.0, n) * 2.0 ** -21              # directions normalised with a 2-ulp reciprocal
    d = [dd[k].astype(F) for k in range(3)]
    cc = sum(c[k].astype(np.float64) ** 2 for k in range(3))
    rr = r.astype(np.float64) ** 2
    w = ((cc - rr) - MARGIN * (cc + rr)).astype(F)
    disc = exact_disc(c, r, o, d)
    val, bound, oo = cheap_value_and_bound(c, w, o, d)
    accepted = ~(disc < 0)
    candidate = ~(val < bound)
    assert accepted.sum() > n // 10 and (~accepted).sum() > n // 10, "the sample straddles the silhouette"
    assert not np.any(accepted & ~candidate), "the conservative mask dropped a primitive the exact test accepts"
    # slack: how much of the margin the disagreement of the two forms uses on accepted primitives
    margin = MARGIN * (oo.astype(np.float64) + cc + rr)
    exact_w = (cc - rr).astype(F).astype(np.float64) - w.astype(np.float64)            # what the table subtracted
    used = ((oo.astype(np.float64) - (val.astype(np.float64) - exact_w)))[accepted] / margin[accepted]   # <= 1 keeps the candidate
    print(f"extent {extent}: largest share of the per-primitive margin used = {used.max():.3f}")
    assert used.max() < 0.5, f"margin nearly exhausted: {used.max():.3f}"
