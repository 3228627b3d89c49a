"""Minimal C3 run for ncu: build the level-9 scene and render a few low-spp frames with the path-loop pipeline."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
from cuda_pathtracer_b200 import scenegen
level = int(sys.argv[1]) if len(sys.argv) > 1 else 9
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
rng = int(sys.argv[3]) if len(sys.argv) > 3 else 0
prims = pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
tris, mat, mats = scenegen.c3_scene(prims, level)
with pkg.Context(0) as ctx:
    scene = ctx.scene_from_triangles(tris, mat, mats, builder=0)
    cam = pkg.default_camera(1920, 1080)
    film = ctx.create_film(1920, 1080)
    for rep in range(3):
        st = ctx.render(scene, cam, film, spp=spp, max_depth=8, pipeline=int(sys.argv[4]) if len(sys.argv) > 4 else 3, rng_mode=rng)
        print(st["segments"] / st["render_ms"] / 1e3, "Mseg/s", st["kernel_launches"], "launches")
