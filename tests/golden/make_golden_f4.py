"""Golden vectors for the any-hit / visibility row: the REFERENCE'S OWN Scene::occluded (scene.h:132-190)
and render_radiosity body (integrator.h:460-504) run on the host through oracle/_ref/ref_host.
Run in the build container (needs /root/reference):  python tests/golden/make_golden_f4.py"""
import json, os, subprocess, tempfile
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "ref_host")
out = {}
rng = np.random.default_rng(41)
n = 3000
lo, hi = np.array([-3.2, -0.3, -6.1], np.float32), np.array([3.2, 5.8, 1.0], np.float32)
segs = np.concatenate([rng.uniform(lo, hi, (n, 3)), rng.uniform(lo, hi, (n, 3))], 1).astype(np.float32)
segs[:200, 3:] = np.float32([0.0, 5.45, -3.0]) + rng.uniform(-0.4, 0.4, (200, 3)).astype(np.float32)     # towards the light
out["segs"] = segs
with tempfile.TemporaryDirectory() as td:
    segs.tofile(os.path.join(td, "segs.bin"))
    for name, cq in (("cbox", 0), ("cbox_quads", 0), ("cbox_quads", 1)):
        tag = f"{name}_cq{cq}"
        scene = os.path.join(ROOT, "scenes", name + ".obj")
        r = subprocess.run([REF, "occluded", scene, os.path.join(td, "segs.bin"), str(n), os.path.join(td, tag), str(cq)], capture_output=True, text=True, check=True)
        print(r.stdout.strip().splitlines()[-1])
        out[tag + "_occ"] = np.fromfile(os.path.join(td, tag + ".occ.u8"), np.uint8)
        out[tag + "_occ_lin"] = np.fromfile(os.path.join(td, tag + ".occ_lin.u8"), np.uint8)
        nprim = {"cbox_cq0": 32, "cbox_quads_cq0": 16, "cbox_quads_cq1": 32}[tag]
        rad = rng.uniform(0.0, 0.6, (nprim, 3)).astype(np.float32)
        rad.tofile(os.path.join(td, "rad.bin"))
        out[tag + "_rad"] = rad
        for (w, h, spp) in ((48, 36, 3), (64, 40, 1)):
            subprocess.run([REF, "visibility", scene, str(w), str(h), str(spp), os.path.join(td, "rad.bin"), os.path.join(td, tag), str(cq)], capture_output=True, text=True, check=True)
            out[f"{tag}_vis_{w}x{h}_s{spp}"] = np.fromfile(os.path.join(td, tag + ".vis.u8"), np.uint8).reshape(h, w, 3)
            out[f"{tag}_visrng_{w}x{h}_s{spp}"] = np.fromfile(os.path.join(td, tag + ".rng.u32"), np.uint32).reshape(h * w, 6)
np.savez_compressed(os.path.join(HERE, "f4_vectors.npz"), **out)
print({k: v.shape for k, v in out.items()})
