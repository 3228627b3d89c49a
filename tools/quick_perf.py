"""Ad-hoc first measurements on the GPU box (not the bench contract; see bench.py)."""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
w, h, spp, depth = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (1920, 1080, 64, 16))]
scene_path = os.path.join(ROOT, "scenes", "cbox_quads.obj")
with pkg.Context(0) as ctx:
    print(ctx.device_info())
    scene = ctx.load_scene(scene_path)
    print(scene.stats())
    cam = pkg.default_camera(w, h)
    t0 = time.time(); film = ctx.create_film(w, h); ctx.sync(); print("film create + rng init s", time.time() - t0)
    for name, kw in [("loop/xorwow", dict(pipeline=2)), ("loop/xorwow/mb4", dict(pipeline=2, loop_min_blocks=4)), ("loop/xorwow/256", dict(pipeline=2, loop_iters=256)), ("loop/philox/mb4", dict(pipeline=2, rng_mode=1, loop_min_blocks=4)), ("loop/xorwow/timed", dict(pipeline=2, stage_timing=True)), ("loop/philox", dict(pipeline=2, rng_mode=1)), ("wavefront/xorwow", {}), ("fused/xorwow", dict(pipeline=1)), ("wavefront/philox", dict(rng_mode=1)),
                     ("fused/philox", dict(rng_mode=1, pipeline=1)), ("wavefront/xorwow/timed", dict(stage_timing=True)),
                     ("wavefront/xorwow/count", dict(count_work=True))]:
        for rep in range(2):
            st = ctx.render(scene, cam, film, spp=spp, max_depth=depth, **kw)
        st["name"] = name
        st["Mseg_per_s"] = st["segments"] / st["render_ms"] / 1e3
        print(json.dumps(st))
ref = os.path.join(ROOT, "oracle", "_ref", "ref_cuda")
for orig in (0, 1):
    r = subprocess.run([ref, "render", scene_path, str(w), str(h), str(spp), str(depth), "2", "-", "0", "0", str(orig)], capture_output=True, text=True)
    print(r.stdout.strip(), r.stderr.strip()[-300:])
