"""Golden vectors on a scene with a real tree (2 566 primitives: the Cornell enclosure + two displaced icospheres of 1 280
triangles, written as .obj by scenegen.write_tree_obj): the REFERENCE'S OWN loadOBJ, BVHBuilder, Scene::intersect (BVH2) and
Scene::intersect_linear, and its integrator, run through oracle/_ref/ref_host on
  * the primitive dump (loader parity on a generated file without `vn`),
  * primary hit ids / distances at 160 x 160,
  * 6 000 adversarial rays (tests/fuzz_scenes.py: aimed at edges, vertices, grazing, started on surfaces),
  * a fixed-seed 48 x 48, 4 spp, depth 8 render (linear radiance).
Run in the build container (needs /root/reference):  python tests/golden/make_golden_tree.py"""
import json, os, subprocess, sys, tempfile
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.dirname(HERE))
import __graft_entry__ as ge
import fuzz_scenes
pkg = ge.load_package()
from cuda_pathtracer_b200 import scenegen
REF = os.path.join(ROOT, "oracle", "_ref", "ref_host")
LEVEL = 3
out = {}
with tempfile.TemporaryDirectory() as td:
    obj = os.path.join(td, "tree.obj")
    base = pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
    n = scenegen.write_tree_obj(obj, base, level=LEVEL)
    run = lambda *a: json.loads(subprocess.run([REF] + [str(x) for x in a], capture_output=True, text=True, check=True).stdout.strip().splitlines()[-1])
    print(run("dump", obj, os.path.join(td, "d"), 0, 0))
    out["prims"] = np.fromfile(os.path.join(td, "d.prims.bin"), np.float32).reshape(-1, 22)
    assert len(out["prims"]) == n
    print(run("ids", obj, 160, 160, os.path.join(td, "i"), 0))
    out["ids"] = np.fromfile(os.path.join(td, "i.ids.i32"), np.int32).reshape(160, 160)
    out["ids_t"] = np.fromfile(os.path.join(td, "i.t.f32"), np.float32).reshape(160, 160)
    prims = pkg.parse_scene_file(obj)
    rays = fuzz_scenes.adversarial_rays(np.random.default_rng(4242), prims, np.zeros(3, np.float32), 3.0, 6000)
    rays.tofile(os.path.join(td, "rays.bin"))
    print(run("rays", obj, os.path.join(td, "rays.bin"), len(rays), os.path.join(td, "r"), 0))
    out["rays"] = rays
    for k, dt in (("bvh", np.int32), ("bvh_t", np.float32), ("lin", np.int32), ("lin_t", np.float32)):
        out["rays_" + k] = np.fromfile(os.path.join(td, "r." + k + (".i32" if dt == np.int32 else ".f32")), dt)
    print(run("render", obj, 48, 48, 4, 8, 1, os.path.join(td, "f"), 0, 1))
    out["render_48x48_4spp_d8"] = np.fromfile(os.path.join(td, "f.linear.f32"), np.float32).reshape(48, 48, 3)
out["level"] = np.int32(LEVEL)
np.savez_compressed(os.path.join(HERE, "tree_vectors.npz"), **out)
print({k: getattr(v, "shape", v) for k, v in out.items()}, os.path.getsize(os.path.join(HERE, "tree_vectors.npz")), "bytes")
