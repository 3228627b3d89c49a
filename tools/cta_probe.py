"""Probe of the CTA-wavefront pipeline (5) against the persistent kernel (3): same image? how fast?
usage: cta_probe.py [scene=c2|c1|c3] [w h spp depth] [variants...]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
which = sys.argv[1] if len(sys.argv) > 1 else "c2"
w, h, spp, depth = [int(x) for x in (sys.argv[2:6] if len(sys.argv) > 5 else (1920, 1080, 64, 16))]
variants = [int(x) for x in sys.argv[6:]] or [0]
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
    out = {}
    for rng in (0, 1):
        base = None
        for name, kw in [("persist", dict(pipeline=3))] + [(f"cta/v{v}", dict(pipeline=5, loop_min_blocks=v)) for v in variants]:
            best = None
            for rep in range(3):
                film.reset()
                st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, rng_mode=rng, **kw)
                best = st if best is None or st["render_ms"] < best["render_ms"] else best
            lin = film.linear()
            if base is None: base = (lin, best["segments"])
            diff = float(np.abs(lin - base[0]).max()); nbad = int((np.abs(lin - base[0]).max(axis=-1) > 0).sum())
            print(json.dumps(dict(name=name, rng=rng, ms=best["render_ms"], segments=best["segments"], paths=best.get("paths"),
                                  Gseg_s=best["segments"] / best["render_ms"] / 1e6, seg_equal=best["segments"] == base[1],
                                  max_abs_diff=diff, pixels_differing=nbad, mean=float(lin.mean()))), flush=True)
