"""Per-pixel mean path length of C2 on the CPU (host build of the product's flat-kernel core, tests/_build/libptx_hostemu.so): a 48 x 27 grid of
sampled pixels, 256 spp each, through a one-pixel sub-window camera.  Output: profiles/r2aa_c2_cost_map.txt.  python tools/cost_map_cpu.py"""
import ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
ge.build_host_emu(); pkg = ge.load_package()
L = C.CDLL(os.path.join(ROOT, "tests", "_build", "libptx_hostemu.so"))
class CamView(C.Structure):
    _fields_ = [("origin", C.c_float*3), ("llc", C.c_float*3), ("horizontal", C.c_float*3), ("vertical", C.c_float*3), ("width", C.c_int), ("height", C.c_int)]
L.emu_scene_load.restype = C.c_void_p; L.emu_scene_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int]
L.emu_render_flat.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
s = L.emu_scene_load(os.path.join(ROOT, "scenes", "cbox_quads.obj").encode(), 0, 0, 3)
W, H = 1920, 1080
c = pkg.default_camera(W, H)
o = np.array(c.origin[:], np.float64); llc = np.array(c.lower_left_corner[:], np.float64); hor = np.array(c.horizontal[:], np.float64); ver = np.array(c.vertical[:], np.float64)
spp = 256
gx, gy = 48, 27
costs = np.zeros((gy, gx))
t = time.time()
for j in range(gy):
    for i in range(gx):
        x0, y0 = int((i + 0.5) * W / gx), int((j + 0.5) * H / gy)
        v = CamView()
        v.origin[:] = o; v.llc[:] = llc + hor * (x0 / W) + ver * (y0 / H); v.horizontal[:] = hor / W; v.vertical[:] = ver / H; v.width = 1; v.height = 1
        lin = np.zeros(3, np.float32); seg = C.c_uint64(); c3 = np.zeros(3, np.uint64)
        L.emu_render_flat(s, C.byref(v), spp, 16, 2023 + j * gx + i, lin.ctypes.data_as(C.c_void_p), None, C.byref(seg), c3.ctypes.data_as(C.c_void_p))
        costs[j, i] = seg.value / spp
print("time", time.time() - t)
np.save('/tmp/costmap.npy', costs)
f = costs.ravel()
print("mean segs/path", f.mean(), "max", f.max(), "quantiles 50/90/95/99", np.quantile(f, [.5, .9, .95, .99]))
for thr in (4, 4.5, 5, 5.5, 6, 6.5, 7, 7.5):
    print("fraction of pixels with > %.1f segs/path: %.4f   share of all work: %.4f" % (thr, (f > thr).mean(), f[f > thr].sum() / f.sum()))
np.set_printoptions(linewidth=250, precision=1)
print(costs[::3, ::3])
