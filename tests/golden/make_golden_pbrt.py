"""Golden vectors for the .pbrt loader: run the REFERENCE's own loadPBRT (utils/pbrt_loader.h, linked with its vendored
pbrtParser) through oracle/_ref/ref_pbrt on the fixtures under tests/pbrt/ and store the primitives it returns.
Run in the build container (needs /root/reference):  python tests/golden/make_golden_pbrt.py"""
import os, subprocess, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
exe = os.path.join(ROOT, "oracle", "_ref", "ref_pbrt")
out = {}
with tempfile.TemporaryDirectory() as td:
    for name in ("basic", "materials", "transforms", "plymesh"):
        subprocess.run([exe, "dump", os.path.join(ROOT, "tests", "pbrt", name + ".pbrt"), os.path.join(td, name)], check=True, capture_output=True)
        r = np.fromfile(os.path.join(td, name + ".prims.bin"), np.float32).reshape(-1, 22)
        out[name + "_v"] = r[:, 1:10].copy(); out[name + "_normal"] = r[:, 13:16].copy()
        out[name + "_bsdf"] = r[:, 16:19].copy(); out[name + "_Le"] = r[:, 19:22].copy()
        out[name + "_type"] = r[:, 0].copy().view(np.int32)
    # a generated, instanced scene of the C5 shape (scenegen.write_c5_pbrt: two object definitions, 27 instances under
    # Translate / Rotate / Scale, emissive quads): too big to store whole, so keep every 61st primitive and a SHA-256 of each array
    import hashlib, sys
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    ge.load_package()
    from cuda_pathtracer_b200 import scenegen
    path = os.path.join(td, "c5_small.pbrt")
    scenegen.write_c5_pbrt(path, instances=27, level=3, n_lights=4)
    subprocess.run([exe, "dump", path, os.path.join(td, "c5_small")], check=True, capture_output=True)
    r = np.fromfile(os.path.join(td, "c5_small.prims.bin"), np.float32).reshape(-1, 22)
    cols = {"v": slice(1, 10), "normal": slice(13, 16), "bsdf": slice(16, 19), "Le": slice(19, 22)}
    out["c5_small_count"] = np.array([len(r)])
    for k, sl in cols.items():
        a = np.ascontiguousarray(r[:, sl])
        out["c5_small_" + k + "_sampled"] = a[::61].copy()
        out["c5_small_" + k + "_sha256"] = np.frombuffer(hashlib.sha256(a.tobytes()).digest(), np.uint8).copy()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "pbrt_vectors.npz"), **out)
print({k: v.shape for k, v in out.items()})
