#!/usr/bin/env python3
"""Generate the golden fixtures from the UNMODIFIED reference (oracle/_ref/ref_driver, built from
/root/reference by oracle/Makefile).  Run in the authoring container only; the .npz files are committed.

Each fixture holds the scene JSON (a tiny-resolution variant of a scenes/*.json file), the reference's
float64 framebuffer and primary-hit id/t buffers for Standard and Paper mode, and the reference's answers
for a batch of seeded random rays (hit, t, p, n, front_face, code, Tracer::trace colour, occlusion probe).
"""
import json, os, sys, tempfile
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_api as oa

CASES = {  # fixture name -> (scene file, dpi)
    "quirks": ("quirks.json", 24), "pokeballs": ("pokeballs.json", 24), "penguin": ("penguin.json", 24),
    "snorlax_reflect": ("snorlax_4k_reflect.json", 18), "synthetic_a": ("synthetic_a_tiny.json", 18),
    "synthetic_b": ("synthetic_b_small.json", 16),
    # the same quirks scene lit by two DirectionalLights as well (scene.h:10-19; pushed into scene.dir_lights by ref_driver --dirlight)
    "quirks_dirlight": ("quirks.json", 24),
}
DIR_LIGHTS = {"quirks_dirlight": [((-0.3, -1.0, -0.5), (0.8, 0.7, 0.6)), ((0.5, -0.2, -1.0), (0.3, 0.3, 0.5))]}

def rays_for(scene: dict, n: int, seed: int) -> np.ndarray:
    """Camera-ish rays plus rays started near/inside the objects' region (exercise secondary-ray quirks)."""
    rng = np.random.default_rng(seed)
    eye = np.array(scene["screen"]["observer"], float)
    P = np.array(scene["screen"]["position"], float); Lx, Ly = scene["screen"]["dimensions"]
    a = np.zeros((n, 6))
    k = n // 2
    S = P + np.stack([rng.uniform(0, Lx, k), rng.uniform(0, Ly, k), np.zeros(k)], 1)
    a[:k, :3] = eye; a[:k, 3:] = S - eye
    centre = P + np.array([Lx / 2, Ly / 2, 0]) + (P + np.array([Lx / 2, Ly / 2, 0]) - eye) * 0.5
    a[k:, :3] = centre + rng.normal(size=(n - k, 3)) * 1.5
    a[k:, 3:] = rng.normal(size=(n - k, 3))
    return a

def main():
    assert oa.have_ref(), "oracle/_ref/ref_driver missing: run `make -C oracle ref` where /root/reference exists"
    only = sys.argv[1:]
    for name, (fname, dpi) in CASES.items():
        if only and name not in only:
            continue
        scene = json.load(open(os.path.join(ROOT, "scenes", fname)))
        scene["screen"]["dpi"] = dpi
        text = json.dumps(scene, separators=(",", ":"))
        with tempfile.NamedTemporaryFile("w", suffix=".json", delete=False) as f:
            f.write(text); path = f.name
        dl = DIR_LIGHTS.get(name, [])
        std = oa.ref_render(path, "standard", want_ids=True, dir_lights=dl)
        pap = oa.ref_render(path, "paper", want_ids=True, dir_lights=dl)
        rays = rays_for(scene, 768, 1234 + len(name))
        rec = oa.ref_trace_rays(path, rays, dir_lights=dl)
        os.unlink(path)
        extra = {"dir_lights": np.array([list(d) + list(r) for d, r in dl], float)} if dl else {}
        np.savez_compressed(os.path.join(HERE, name + ".npz"), scene_json=np.array(text), **extra,
                            std_fb=std["fb"], std_ids=std["ids"], std_t=std["t"],
                            paper_fb=pap["fb"], paper_ids=pap["ids"], paper_t=pap["t"], rays=rays, ray_records=rec)
        print(name, std["fb"].shape, os.path.getsize(os.path.join(HERE, name + ".npz")) // 1024, "KiB")

if __name__ == "__main__":
    main()
