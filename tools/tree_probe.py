"""CPU-side look at a BVH8 (host builder + the product's traversal compiled for the host, tests/_build/libptx_hostemu.so):
node layout and node visits / triangle tests per path segment over a real path-traced ray distribution.
    python tools/tree_probe.py [scene.obj] [max_leaf] [w h spp depth]"""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
ge.build()
pkg = ge.load_package()
L = C.CDLL(os.path.join(ROOT, "tests", "_build", "libptx_hostemu.so"))
L.emu_scene_load.restype = C.c_void_p
L.emu_scene_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int]
L.emu_scene_nodes.restype = C.c_int64
L.emu_scene_nodes.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
L.emu_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
L.emu_render_counters.argtypes = [C.c_void_p]
L.emu_scene_counts.restype = C.c_int64
L.emu_scene_counts.argtypes = [C.c_void_p] * 5
path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "scenes", "cbox_quads.obj")
max_leaf = int(sys.argv[2]) if len(sys.argv) > 2 else 3
w, h, spp, depth = [int(x) for x in sys.argv[3:7]] if len(sys.argv) > 6 else (96, 54, 8, 16)
s = L.emu_scene_load(path.encode(), 0, 0, max_leaf)
nt, nn, dp, sah = C.c_int64(), C.c_int64(), C.c_int(), C.c_float()
L.emu_scene_counts(s, C.byref(nt), C.byref(nn), C.byref(dp), C.byref(sah))
print(f"tris {nt.value} nodes {nn.value} depth {dp.value} sah {sah.value:.4f}")
nodes = np.zeros((nn.value, 20), np.uint32)
L.emu_scene_nodes(s, nodes.ctypes.data_as(C.c_void_p), nn.value)
if nn.value <= 64:
    for i, n in enumerate(nodes):
        p = n[:3].view(np.float32); e = [(n[3] >> (8 * k)) & 255 for k in range(3)]; imask = n[3] >> 24
        meta = np.array([n[6], n[7]], np.uint32).view(np.uint8)
        q = n[8:20].view(np.uint8).reshape(6, 8)
        print(f"node {i}: p {p} e {e} imask {imask:08b} child_base {n[4]} tri_base {n[5]}")
        for sl in range(8):
            if meta[sl] == 0: continue
            sc = [0.0 if e[k] == 0 else 2.0 ** (e[k] - 127) for k in range(3)]
            lo = [p[k] + q[k, sl] * sc[k] for k in range(3)]; hi = [p[k] + q[3 + k, sl] * sc[k] for k in range(3)]
            kind = "inner" if (imask >> sl) & 1 else f"leaf ntri {bin(meta[sl] >> 5).count('1')} off {meta[sl] & 31}"
            print(f"   slot {sl}: {kind:22s} lo {np.round(lo, 3)} hi {np.round(hi, 3)}")
class CamView(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("llc", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("width", C.c_int), ("height", C.c_int)]
c = pkg.default_camera(w, h)
v = CamView()
v.origin[:] = c.origin; v.llc[:] = c.lower_left_corner; v.horizontal[:] = c.horizontal; v.vertical[:] = c.vertical
v.width, v.height = w, h
seg = C.c_uint64()
L.emu_render(s, C.byref(v), spp, depth, 2023, None, None, C.byref(seg))
ctr = (C.c_uint64 * 2)()
L.emu_render_counters(ctr)
print(f"segments {seg.value}  node visits/segment {ctr[0] / seg.value:.3f}  triangle tests/segment {ctr[1] / seg.value:.3f}")
