"""C3 probe on the GPU box: generate the displaced-icosphere scene, build on the device, measure traversal work and throughput."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
from cuda_pathtracer_b200 import scenegen
level = int(sys.argv[1]) if len(sys.argv) > 1 else 9
w, h, spp, depth = 1920, 1080, int(sys.argv[2]) if len(sys.argv) > 2 else 16, 8
prims = pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
t0 = time.time(); tris, mat, mats = scenegen.c3_scene(prims, level); print("generate s", round(time.time() - t0, 2), tris.shape, flush=True)
with pkg.Context(0) as ctx:
    t0 = time.time(); scene = ctx.scene_from_triangles(tris, mat, mats, builder=0); print("scene_from_triangles s", round(time.time() - t0, 2))
    print(json.dumps(scene.stats()), flush=True)
    cam = pkg.default_camera(w, h)
    film = ctx.create_film(w, h)
    for name, kw in [("loop/xorwow", dict(pipeline=2)), ("loop/xorwow/mb4", dict(pipeline=2, loop_min_blocks=4)), ("loop/philox/mb4", dict(pipeline=2, rng_mode=1, loop_min_blocks=4)), ("loop/philox", dict(pipeline=2, rng_mode=1)), ("fused/xorwow", dict(pipeline=1)),
                     ("wavefront/xorwow", dict(pipeline=0)), ("wavefront/xorwow/timed", dict(pipeline=0, stage_timing=True)),
                     ("wavefront/xorwow/count", dict(pipeline=0, count_work=True))]:
        for rep in range(2):
            st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, **kw)
        st["name"] = name; st["Mseg_per_s"] = st["segments"] / st["render_ms"] / 1e3
        if st["node_visits"]:
            st["nodes_per_seg"] = st["node_visits"] / st["segments"]; st["tris_per_seg"] = st["tri_tests"] / st["segments"]
        print(json.dumps(st), flush=True)
    film.write(os.path.join(ROOT, "gpurun_out", f"c3_level{level}.ppm"))
    if len(sys.argv) > 3:
        t0 = time.time(); scene.rebuild(builder=1); print("host rebuild s", round(time.time() - t0, 2)); print(json.dumps(scene.stats()))
        st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, pipeline=2); print("host-built loop Mseg/s", st["segments"] / st["render_ms"] / 1e3)
