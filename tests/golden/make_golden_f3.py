"""Golden vectors for geometry preprocessing: the REFERENCE'S OWN convertQuadsToTriangles
(application_state.h:323-365, restated in the harness) followed by its subdivide_primitives (form_factors.h:475-574),
in loadScene's order (application_state.h:403-412), dumped through oracle/_ref/ref_host.
Run in the build container (needs /root/reference):  python tests/golden/make_golden_f3.py"""
import os, subprocess, tempfile
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "ref_host")
out = {}
with tempfile.TemporaryDirectory() as td:
    for name, cq, sub in (("cbox", 0, 1), ("cbox", 0, 3), ("cbox_quads", 0, 2), ("cbox_quads", 1, 1), ("cbox_quads", 1, 2)):
        tag = f"{name}_cq{cq}_sub{sub}"
        subprocess.run([REF, "dump", os.path.join(ROOT, "scenes", name + ".obj"), os.path.join(td, tag), str(cq), str(sub)], capture_output=True, text=True, check=True)
        out[tag] = np.fromfile(os.path.join(td, tag + ".prims.bin"), np.float32).reshape(-1, 22)
np.savez_compressed(os.path.join(HERE, "f3_vectors.npz"), **out)
print({k: v.shape for k, v in out.items()})
