"""One scene, a few frames of one pipeline: the process ncu attaches to.
    python tools/ncu_case.py c1|c2|c4|c3[:level] w h spp depth pipeline rng min_blocks [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
which = sys.argv[1]
w, h, spp, depth, pipe, rng, minb = [int(x) for x in sys.argv[2:9]]
frames = int(sys.argv[9]) if len(sys.argv) > 9 else 2
with pkg.Context(0) as ctx:
    cam = None
    if which == "c5":                     # SURVEY 8(d) C5 as bench.py builds it: 305 instances x 327 680 triangles, two-level
        from cuda_pathtracer_b200 import scenegen
        scenegen.write_c5_pbrt("/tmp/c5_ncu.pbrt", instances=305, level=7)
        scene = ctx.load_scene("/tmp/c5_ncu.pbrt", mirror_from_mtl=True, keep_instances=True)
        room = 28.0
        cam = pkg.look_at_camera((room / 2, room / 2, room * 1.9), (room / 2, room / 2, 0.0), (0.0, 1.0, 0.0), 40.0, w, h)
    elif which.startswith("c3") or which == "c4":
        from cuda_pathtracer_b200 import scenegen
        base = pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
        if which == "c4":
            scene = ctx.scene_from_primitives(scenegen.c4_prims(base), builder=0)
        else:
            tris, mat, mats = scenegen.c3_scene(base, int(which.split(":")[1]) if ":" in which else 9)
            scene = ctx.scene_from_triangles(tris, mat, mats, builder=0)
    else:
        scene = ctx.load_scene(os.path.join(ROOT, "scenes", "cbox.obj" if which == "c1" else "cbox_quads.obj"))
    cam = cam or pkg.default_camera(w, h)
    film = ctx.create_film(w, h)
    for rep in range(frames):
        st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, pipeline=pipe, rng_mode=rng, loop_min_blocks=minb)
        print(which, "pipeline", pipe, "rng", rng, "minb", minb, round(st["segments"] / st["render_ms"] / 1e3), "Mseg/s", st["segments"], "segments", round(st["render_ms"], 3), "ms", flush=True)
