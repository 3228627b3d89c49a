"""Persistent-kernel throughput probe (best of 3 frames): perf_probe.py c2|c3|c1 w h spp depth [rng...].  PTX_LIB selects the library."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
which = sys.argv[1]
w, h, spp, depth = [int(x) for x in sys.argv[2:6]]
rngs = [int(x) for x in sys.argv[6:]] or [0]
with pkg.Context(0) as ctx:
    if which == "c3":
        from cuda_pathtracer_b200 import scenegen
        base = pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
        tris, mat, mats = scenegen.c3_scene(base, int(os.environ.get("C3_LEVEL", "9")))
        scene = ctx.scene_from_triangles(tris, mat, mats, builder=0)
    else:
        scene = ctx.load_scene(os.path.join(ROOT, "scenes", "cbox.obj" if which == "c1" else "cbox_quads.obj"))
    cam = pkg.default_camera(w, h)
    film = ctx.create_film(w, h)
    for rng in rngs:
        best = None
        for rep in range(4):
            film.reset()
            st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, rng_mode=rng, loop_min_blocks=int(os.environ.get("MINB", "0")), shade_batch=int(os.environ.get("BATCH", "0")), chunks=int(os.environ.get("CHUNKS", "0")))
            best = st if best is None or st["render_ms"] < best["render_ms"] else best
        print(json.dumps(dict(minb=os.environ.get("MINB", "0"), batch=os.environ.get("BATCH", "0"), lib=os.path.basename(pkg.LIB_PATH), scene=which, rng=rng, ms=round(best["render_ms"], 3), segments=best["segments"],
                              Gseg_s=round(best["segments"] / best["render_ms"] / 1e6, 3), mean=float(film.linear().mean()))), flush=True)
