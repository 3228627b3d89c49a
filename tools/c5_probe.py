"""C5-shaped scene end to end on the box: write the instanced .pbrt, load it through ptx_scene_load (parse, flatten the
instances, pack, device BVH build) and render a few frames.   python tools/c5_probe.py [instances] [level] [w h spp]"""
import os, sys, time, json, resource
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
from cuda_pathtracer_b200 import scenegen
inst = int(sys.argv[1]) if len(sys.argv) > 1 else 4883
level = int(sys.argv[2]) if len(sys.argv) > 2 else 5
w, h, spp = [int(x) for x in sys.argv[3:6]] if len(sys.argv) > 5 else (1920, 1080, 16)
batches = [int(x) for x in sys.argv[6].split(',')] if len(sys.argv) > 6 else [0]
path = "/tmp/c5.pbrt"
t = time.time(); n = scenegen.write_c5_pbrt(path, inst, level); t_write = time.time() - t
with pkg.Context(0) as ctx:
    t = time.time(); scene = ctx.load_scene(path, mirror_from_mtl=True); t_load = time.time() - t
    st = scene.stats()
    side = int(-(-inst ** (1 / 3.0) // 1)); room = 4.0 * side
    cam = pkg.look_at_camera((room / 2, room / 2, room * 1.9), (room / 2, room / 2, 0), (0, 1, 0), 40.0, w, h)
    film = ctx.create_film(w, h)
    out = {"triangles": n, "write_s": round(t_write, 2), "load_s": round(t_load, 2), "build_ms": st["build_ms"], "nodes": st["n_nodes"], "sah": st["sah_cost"],
           "max_depth": st["max_depth"], "node_GB": st["node_bytes"] / 1e9, "tri_GB": st["triangle_bytes"] / 1e9,
           "host_peak_GB": resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6}
    for sb in batches:
      for rng in (0, 1):
        for rep in range(2):
            r = ctx.render(scene, cam, film, spp=spp, max_depth=8, rng_mode=rng, shade_batch=sb, loop_min_blocks=int(os.environ.get("MINB", "0")))
        out[f"Gseg_s_rng{rng}_sb{sb}"] = round(r["segments"] / r["render_ms"] / 1e6, 3); out[f"seg_per_path_rng{rng}"] = r["segments"] / r["paths"]
    film.write(os.path.join(ROOT, "gpurun_out", "c5.png"))
    print(json.dumps(out))
