"""Round-2 probe (not the bench contract): flat-list kernel against the persistent BVH8 kernel and the reference's CUDA kernels.
    python tools/r2_probe.py [spp]
Prints one JSON line per measurement; compares the films of the two pipelines bit for bit."""
import json, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 256
cases = [("c2", "cbox_quads.obj", 1920, 1080, spp, 16), ("c1", "cbox.obj", 512, 512, 64, 8)]
with pkg.Context(0) as ctx:
    for name, scene_file, w, h, s, depth in cases:
        scene = ctx.load_scene(os.path.join(ROOT, "scenes", scene_file))
        print(json.dumps({"scene": name, "stats": {k: v for k, v in scene.stats().items() if k in ("n_triangles", "n_nodes", "flat_units", "build_ms")}}), flush=True)
        cam = pkg.default_camera(w, h)
        film = ctx.create_film(w, h)
        films = {}
        for rng in (0, 1):
            for label, kw in (("persist", dict(pipeline=3)), ("flat/3", dict(pipeline=6, loop_min_blocks=3)), ("flat/4", dict(pipeline=6, loop_min_blocks=4)), ("default", dict())):
                best = None
                for rep in range(3):
                    film.reset()
                    st = ctx.render(scene, cam, film, spp=s, max_depth=depth, rng_mode=rng, **kw)
                    best = st if best is None or st["render_ms"] < best["render_ms"] else best
                lin = film.linear()
                films[(rng, label)] = lin
                print(json.dumps(dict(scene=name, rng=rng, pipeline=label, ms=round(best["render_ms"], 3), segments=best["segments"],
                                      Gseg_s=round(best["segments"] / best["render_ms"] / 1e6, 3), mean=float(lin.mean()))), flush=True)
            a, b = films[(rng, "persist")], films[(rng, "flat/3")]
            diff = (a != b).any(-1)
            print(json.dumps(dict(scene=name, rng=rng, check="flat vs persist", differing_pixels=int(diff.sum()), of=int(diff.size),
                                  max_abs=float(np.abs(a - b).max()), flat4_equals_flat3=bool(np.array_equal(films[(rng, "flat/4")], b)))), flush=True)
        film.close(); scene.close()
ref = os.path.join(ROOT, "oracle", "_ref", "ref_cuda")
for name, scene_file, w, h, s, depth in cases:
    for orig in (0, 1):
        r = subprocess.run([ref, "render", os.path.join(ROOT, "scenes", scene_file), str(w), str(h), str(s), str(depth), "2", "-", "0", "0", str(orig)], capture_output=True, text=True)
        print(r.stdout.strip(), r.stderr.strip()[-300:], flush=True)
