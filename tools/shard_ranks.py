"""Per-rank cost of a tile-sharded frame, measured on ONE GPU: renders rank r of `world` for every r and prints ms / segments.
shard_ranks.py w h spp depth world [chunks...]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
w, h, spp, depth, world = [int(x) for x in sys.argv[1:6]]
chunk_list = [int(x) for x in sys.argv[6:]] or [0]
with pkg.Context(0) as ctx:
    scene = ctx.load_scene(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
    cam = pkg.default_camera(w, h)
    film = ctx.create_film(w, h)
    for chunks in chunk_list:
        rows = []
        for r in (range(1) if os.environ.get('ONLY_RANK0') else range(world)):
            best = None
            for rep in range(3):
                film.reset()
                st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, rng_mode=0, shard_mode=pkg.SHARD_TILES if world > 1 else pkg.SHARD_NONE,
                                shard_rank=r, shard_world=world, chunks=chunks, loop_min_blocks=int(os.environ.get('MINB', '0')))
                best = st if best is None or st["render_ms"] < best["render_ms"] else best
            rows.append((round(best["render_ms"], 2), best["segments"]))
        tot = sum(s for _, s in rows)
        print(json.dumps(dict(chunks=chunks, world=world, ms=[m for m, _ in rows], Mseg=[round(s / 1e6, 1) for _, s in rows],
                              max_ms=max(m for m, _ in rows), agg_Gseg_s=round(tot / max(m for m, _ in rows) / 1e6, 2))), flush=True)
