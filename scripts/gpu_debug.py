import importlib, os, sys, collections
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
rt = importlib.import_module("raytracing-project_b200")
import oracle_api as oa
name, mode = sys.argv[1], sys.argv[2]
path = os.path.join(ROOT, "scenes", name + ".json")
rt.init(0)
sc = rt.load_scene_from_json(path)
tr = rt.Tracer(sc, rt.RenderMode.Paper if mode == "paper" else rt.RenderMode.Standard)
g = tr.render(want_rgb8=True, want_ids=True)
orc = oa.Oracle.from_file(path)
o = orc.render(mode, want_ids=True)
ob = oa.to_byte(o["fb"])
diff = np.abs(g["rgb8"].astype(int) - ob.astype(int)).max(axis=2)
bad = diff > 1
print("bad pixels", bad.sum(), "of", bad.size)
ids0 = o["ids"][..., 0] // 8
print("bad by primary object (oracle sample 0):", collections.Counter(ids0[bad].tolist()).most_common(12))
print("all by primary object:", collections.Counter(ids0.ravel().tolist()).most_common(12))
ys, xs = np.nonzero(bad)
for y, x in list(zip(ys, xs))[:12]:
    print((x, y), "gpu", g["rgb8"][y, x], "orc", ob[y, x], "ids gpu", g["ids"][y, x], "orc", o["ids"][y, x])
# per-ray comparison on the centre rays of some bad pixels
import json
scj = json.load(open(path)); scr = scj["screen"]
eye = np.array(scr["observer"], float); P = np.array(scr["position"], float); Lx, Ly = scr["dimensions"]
W, H = sc.width, sc.height
rays = []
for y, x in list(zip(ys, xs))[:2000]:
    sx, sy = (x + 0.5) / W, (H - 1 - y + 0.5) / H
    S = P + np.array([Lx * sx, Ly * sy, 0.0]); d = S - eye; d /= np.linalg.norm(d)
    rays.append(np.concatenate([eye, d]))
if rays:
    rays = np.array(rays)
    a = tr.trace_rays(rays); b = orc.trace_rays(rays)
    cd = np.abs(a[:, 10:13] - b[:, 10:13]).max(axis=1)
    print("centre rays: color diff > 0.01:", (cd > 0.01).sum(), "of", len(rays))
    k = np.argsort(-cd)[:8]
    np.set_printoptions(precision=5, suppress=True, linewidth=200)
    for i in k:
        print("ray", i, "gpu", a[i, :14]); print("      orc", b[i, :14])
