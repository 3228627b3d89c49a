"""Device BVH builder timing: C3 at several subdivision levels.  python tools/build_probe.py [levels...]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
from cuda_pathtracer_b200 import scenegen
levels = [int(x) for x in sys.argv[1:]] or [7, 9]
base = pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
with pkg.Context(0) as ctx:
    for level in levels:
        tris, mat, mats = scenegen.c3_scene(base, level)
        for rep in range(2):
            t0 = time.time()
            scene = ctx.scene_from_triangles(tris, mat, mats, builder=0)
            wall = time.time() - t0
            st = scene.stats()
            scene.close()
        print(json.dumps(dict(level=level, triangles=st["n_triangles"], nodes=st["n_nodes"], build_ms=round(st["build_ms"], 1),
                              stages_ms=[round(x, 1) for x in st["build_stage_ms"]], sah=round(st["sah_cost"], 4), depth=st["max_depth"],
                              wall_s=round(wall, 2))), flush=True)
