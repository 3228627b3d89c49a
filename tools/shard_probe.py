"""How well does ONE rank's share of a frame scale?  Renders rank 0's tiles of an N-way tile-sharded frame on one GPU
(no NCCL) for N = 1, 2, 4, 8 and prints kernel ms against the ideal T1 / N.   python tools/shard_probe.py c2|c3 [chunks]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rng = int(sys.argv[3]) if len(sys.argv) > 3 else 0
min_blocks = int(sys.argv[4]) if len(sys.argv) > 4 else 0
pipe = int(sys.argv[5]) if len(sys.argv) > 5 else 0      # 0 = the library default (flat kernel for the Cornell scenes)
worlds = [int(x) for x in sys.argv[6].split(",")] if len(sys.argv) > 6 else (1, 2, 4, 8)
with pkg.Context(0) as ctx:
    if wl == "c3":
        from cuda_pathtracer_b200 import scenegen
        tris, mat, mats = scenegen.c3_scene(pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj")), 9)
        scene = ctx.scene_from_triangles(tris, mat, mats, builder=0)
        w, h, spp, depth = 1920, 1080, 256, 8
    else:
        scene = ctx.load_scene(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
        w, h, spp, depth = 1920, 1080, 1024, 16
    cam = pkg.default_camera(w, h)
    film = ctx.create_film(w, h)
    t1 = None
    for n in worlds:
        mode = pkg.SHARD_NONE if n == 1 else (pkg.SHARD_TILES if rng == 0 else pkg.SHARD_SAMPLES)
        for rep in range(2):
            st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, pipeline=pipe, cost_order=os.environ.get("COST_ORDER", "1") == "1", chunks=chunks, rng_mode=rng, loop_min_blocks=min_blocks,
                            shard_mode=mode, shard_rank=0, shard_world=n)
        t1 = t1 or st["render_ms"]
        print(json.dumps({"workload": wl, "chunks": chunks, "rng": rng, "min_blocks": min_blocks, "pipeline": pipe, "cost_order": os.environ.get("COST_ORDER", "1"), "world": n, "render_ms": round(st["render_ms"], 2), "ideal_ms": round(t1 / n, 2),
                          "efficiency": round(t1 / n / st["render_ms"], 3), "Gseg_s": round(st["segments"] / st["render_ms"] / 1e6, 2)}))
