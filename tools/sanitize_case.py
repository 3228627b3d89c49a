"""Small renders that exercise every hand-off path of the product kernels; run under compute-sanitizer by tools/sanitize.sh."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
w, h, spp = 48, 27, 64            # 1 296 pixels for ~100 K resident lanes: every chunk waits for its predecessor (the `pending` path)
with pkg.Context(0) as ctx:
    for scene_file in ("cbox_quads.obj", "cbox.obj"):
        scene = ctx.load_scene(os.path.join(ROOT, "scenes", scene_file))
        cam = pkg.default_camera(w, h)
        film = ctx.create_film(w, h)
        for pipe in (pkg.PIPELINE_PERSIST, pkg.PIPELINE_FLAT):
            for rng in (pkg.RNG_XORWOW_COMPAT, pkg.RNG_PHILOX):
                for chunks in (0, -8):
                    film.reset()
                    st = ctx.render(scene, cam, film, spp=spp, max_depth=16, pipeline=pipe, rng_mode=rng, chunks=chunks)
                    film.rgb8()
                    print(scene_file, "pipeline", pipe, "rng", rng, "chunks", chunks, "segments", st["segments"], flush=True)
        ids, ts = ctx.trace_primary_ids(scene, cam, w, h)
        film.close(); scene.close()
    # a mid-size scene through the device builder + the two-records-per-round traversal
    prims = pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox.obj"), subdivision=3)
    scene = ctx.scene_from_primitives(prims, builder=0)
    film = ctx.create_film(w, h)
    for rng in (0, 1):
        st = ctx.render(scene, pkg.default_camera(w, h), film, spp=16, max_depth=8, rng_mode=rng)
        print("subdivided", len(prims), "rng", rng, "segments", st["segments"], flush=True)
    film.close(); scene.close()
print("sanitize_case done")
