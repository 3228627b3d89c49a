"""Golden vectors for the grid-guided / MIS sampling modes, made by running the REFERENCE's own
integrator + Grid code on the host (oracle/_ref/ref_host guided ...) in the build container.

Inputs are synthetic: per primitive a 16x16 RGB "radiosity grid" (what the out-of-scope radiosity solver would deliver)
from `synthetic_grids` below -- a smooth lobe plus noise, with whole-zero grids (invalid -> cosine fallback), zero rows
(uniform conditional CDF) and zero cells (pdf floor 1e-6) mixed in so every branch of grid.h is exercised.

    python tests/golden/make_golden_guided.py      # needs /root/reference (build container only)
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "ref_host")


def synthetic_grids(n_prims, seed=4242):
    """[n_prims][256][3] float32; deterministic.  Row r = theta band (rows 0..7 = upper hemisphere), column = phi band."""
    rng = np.random.RandomState(seed)
    g = np.zeros((n_prims, 16, 16, 3), np.float32)
    rows = np.arange(16, dtype=np.float32)[:, None]
    cols = np.arange(16, dtype=np.float32)[None, :]
    for p in range(n_prims):
        r0, c0 = rng.uniform(0, 7), rng.uniform(0, 16)
        dc = np.minimum(np.abs(cols - c0), 16 - np.abs(cols - c0))
        lobe = np.exp(-((rows - r0) ** 2) / 6.0 - (dc ** 2) / 10.0).astype(np.float32)
        base = (0.05 + 0.1 * rng.rand(16, 16)).astype(np.float32)
        lum = (lobe * np.float32(2.0) + base).astype(np.float32)
        tint = rng.uniform(0.5, 1.0, 3).astype(np.float32)
        g[p] = lum[:, :, None] * tint[None, None, :]
        kind = p % 7
        if kind == 3:
            g[p] = 0.0                                  # invalid grid: cosine fallback
        elif kind == 4:
            g[p, 2, :, :] = 0.0; g[p, 5, :, :] = 0.0    # two empty rows: uniform row CDFs, never selected by the marginal
        elif kind == 5:
            g[p, :, ::3, :] = 0.0                       # empty cells: pdf floor
    return g.reshape(n_prims, 256, 3)


CASES = [  # tag, scene, convert_quads, w, h, spp, depth, mode, bsdf_fraction
    ("cbox_mis", "cbox.obj", 0, 40, 32, 8, 6, 3, 0.5),
    ("cbox_grid", "cbox.obj", 0, 40, 32, 8, 6, 2, 0.5),
    ("quads_mis30", "cbox_quads.obj", 0, 36, 28, 8, 8, 3, 0.3),
    ("quads_grid", "cbox_quads.obj", 0, 36, 28, 6, 8, 1, 0.5),
    ("quads_topk", "cbox_quads.obj", 1, 24, 24, 4, 5, 4, 0.5),
]


def main():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    pkg = ge.load_package()
    out = {}
    meta = {}
    with tempfile.TemporaryDirectory() as td:
        for tag, scene, cq, w, h, spp, depth, mode, frac in CASES:
            path = os.path.join(ROOT, "scenes", scene)
            n = len(pkg.parse_scene_file(path, convert_quads=bool(cq)))
            grids = synthetic_grids(n)
            gf = os.path.join(td, tag + ".grids.f32")
            grids.tofile(gf)
            pre = os.path.join(td, tag)
            r = subprocess.run([REF, "guided", path, str(w), str(h), str(spp), str(depth), str(mode), repr(frac), gf, pre, str(cq)],
                               capture_output=True, text=True, check=True)
            info = json.loads(r.stdout.strip().splitlines()[-1])
            out[tag + "_linear"] = np.fromfile(pre + ".linear.f32", np.float32).reshape(h, w, 3)
            out[tag + "_rng"] = np.fromfile(pre + ".rng.u32", np.uint32).reshape(h * w, 6)
            meta[tag] = dict(scene=scene, convert_quads=cq, w=w, h=h, spp=spp, depth=depth, mode=mode, bsdf_fraction=frac, n_prims=n,
                             mean=info["mean"])
            print(tag, info["mean"])
    np.savez_compressed(os.path.join(HERE, "guided_vectors.npz"), **out)
    with open(os.path.join(HERE, "guided_vectors.json"), "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
