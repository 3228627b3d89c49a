"""Golden vectors for the .pbrt loader: run the REFERENCE's own loadPBRT (utils/pbrt_loader.h, linked with its vendored
pbrtParser) through oracle/_ref/ref_pbrt on the fixtures under tests/pbrt/ and store the primitives it returns.
Run in the build container (needs /root/reference):  python tests/golden/make_golden_pbrt.py"""
import os, subprocess, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
exe = os.path.join(ROOT, "oracle", "_ref", "ref_pbrt")
out = {}
with tempfile.TemporaryDirectory() as td:
    for name in ("basic", "materials", "transforms"):
        subprocess.run([exe, "dump", os.path.join(ROOT, "tests", "pbrt", name + ".pbrt"), os.path.join(td, name)], check=True, capture_output=True)
        r = np.fromfile(os.path.join(td, name + ".prims.bin"), np.float32).reshape(-1, 22)
        out[name + "_v"] = r[:, 1:10].copy(); out[name + "_normal"] = r[:, 13:16].copy()
        out[name + "_bsdf"] = r[:, 16:19].copy(); out[name + "_Le"] = r[:, 19:22].copy()
        out[name + "_type"] = r[:, 0].copy().view(np.int32)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "pbrt_vectors.npz"), **out)
print({k: v.shape for k, v in out.items()})
