"""Golden vectors for the END-POINT RULE of the any-hit query: segments whose far end lies within a few ulps of a surface, so that
the hit distance equals t_max = dist - 1e-4 exactly for some of them.  The reference's Triangle::intersect accepts t == t_max
(triangle.h:83), its Quad halves do not (quad.h:76, 106); expected answers come from the REFERENCE'S OWN Scene::occluded run on the
host (oracle/_ref/ref_host occluded).  The oracle only steers the bisection towards the boundary.
Run in the build container (needs /root/reference):  python tests/golden/make_golden_f4_boundary.py"""
import os, subprocess, sys, tempfile
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_binding
REF = os.path.join(ROOT, "oracle", "_ref", "ref_host")
f32 = np.float32
o = oracle_binding.Oracle()
osc = o.load_obj(os.path.join(ROOT, "scenes", "cbox.obj"))
rng = np.random.default_rng(77)
segs = []
lo, hi = np.array([-2.8, 0.1, -5.5], f32), np.array([2.4, 5.0, -0.5], f32)
while len(segs) < 400 * 13:
    a = rng.uniform(lo, hi, 3).astype(f32)
    u = rng.normal(size=3); u = (u / np.linalg.norm(u)).astype(f32)
    seg = lambda s: np.concatenate([a, (a + u * f32(s)).astype(f32)]).astype(f32)
    s_lo, s_hi = f32(0.01), f32(30.0)
    if o.occluded(osc, seg(s_lo), mode=1)[0] or not o.occluded(osc, seg(s_hi), mode=1)[0]:
        continue
    while np.nextafter(s_lo, f32(np.inf), dtype=f32) < s_hi:          # float bisection down to adjacent values
        mid = f32((np.float64(s_lo) + np.float64(s_hi)) / 2)
        if o.occluded(osc, seg(mid), mode=1)[0]: s_hi = mid
        else: s_lo = mid
    s = s_hi
    for k in range(6): s = np.nextafter(s, f32(0), dtype=f32)
    for k in range(13):
        segs.append(seg(s)); s = np.nextafter(s, f32(np.inf), dtype=f32)
segs = np.array(segs, f32)
out = {"segs": segs}
with tempfile.TemporaryDirectory() as td:
    segs.tofile(os.path.join(td, "segs.bin"))
    for name, cq in (("cbox", 0), ("cbox_quads", 0), ("cbox_quads", 1)):
        tag = f"{name}_cq{cq}"
        r = subprocess.run([REF, "occluded", os.path.join(ROOT, "scenes", name + ".obj"), os.path.join(td, "segs.bin"), str(len(segs)), os.path.join(td, tag), str(cq)],
                           capture_output=True, text=True, check=True)
        print(tag, r.stdout.strip().splitlines()[-1])
        out[tag + "_occ"] = np.fromfile(os.path.join(td, tag + ".occ.u8"), np.uint8)
        out[tag + "_occ_lin"] = np.fromfile(os.path.join(td, tag + ".occ_lin.u8"), np.uint8)
print("quads vs the same geometry as triangles differ on", int((out["cbox_quads_cq0_occ_lin"] != out["cbox_quads_cq1_occ_lin"]).sum()), "segments (the t == t_max cases)")
np.savez_compressed(os.path.join(HERE, "f4_boundary.npz"), **out)
