"""Wavefront pipeline with / without queue sorting against the persistent kernel.  python tools/sort_probe.py c3|c2 [spp]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
with pkg.Context(0) as ctx:
    if wl == "c3":
        from cuda_pathtracer_b200 import scenegen
        tris, mat, mats = scenegen.c3_scene(pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj")), 9)
        scene = ctx.scene_from_triangles(tris, mat, mats, builder=0)
        w, h, spp, depth = 1920, 1080, 64, 8
    else:
        scene = ctx.load_scene(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
        w, h, spp, depth = 1920, 1080, 64, 16
    if len(sys.argv) > 2: spp = int(sys.argv[2])
    cam = pkg.default_camera(w, h)
    film = ctx.create_film(w, h)
    for rng in (0, 1):
        for label, kw in (("persistent (default)", dict()), ("wavefront", dict(pipeline=pkg.PIPELINE_WAVEFRONT)),
                          ("wavefront + ray sort", dict(pipeline=pkg.PIPELINE_WAVEFRONT, sort_rays=True)),
                          ("wavefront + material sort", dict(pipeline=pkg.PIPELINE_WAVEFRONT, sort_material=True)),
                          ("wavefront + both", dict(pipeline=pkg.PIPELINE_WAVEFRONT, sort_rays=True, sort_material=True))):
            best = None
            for rep in range(2):
                film.reset()
                st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, rng_mode=rng, **kw)
                best = st if best is None or st["render_ms"] < best["render_ms"] else best
            timed = ctx.render(scene, cam, film, spp=spp, max_depth=depth, rng_mode=rng, stage_timing=True, **kw) if "pipeline" in kw else None
            print(json.dumps(dict(workload=wl, rng=rng, variant=label, ms=round(best["render_ms"], 2), Gseg_s=round(best["segments"] / best["render_ms"] / 1e6, 3),
                                  launches=best["kernel_launches"], extend_ms=round(timed["extend_ms"], 2) if timed else None,
                                  shade_ms=round(timed["shade_ms"], 2) if timed else None)), flush=True)
