"""ctypes binding over oracle/liboracle.so -- TEST INFRASTRUCTURE (checker only, never the product path)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "liboracle.so")
REF_HOST = os.path.join(ROOT, "oracle", "_ref", "ref_host")
REF_CUDA = os.path.join(ROOT, "oracle", "_ref", "ref_cuda")


class XW(C.Structure):
    _fields_ = [("v", C.c_uint32 * 5), ("d", C.c_uint32)]


class Cam(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("llc", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("width", C.c_int), ("height", C.c_int)]


OPRIM_DTYPE = np.dtype([("type", "<i4"), ("v", "<f4", (4, 3)), ("n", "<f4", 3), ("bsdf", "<f4", 3), ("Le", "<f4", 3)])
ONODE_DTYPE = np.dtype([("bmin", "<f4", 3), ("bmax", "<f4", 3), ("left", "<i4"), ("right", "<i4"), ("count", "<i4")])


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self):
        L = C.CDLL(LIB)
        L.pto_uniform.restype = C.c_float
        L.pto_load_obj.restype = C.c_void_p
        L.pto_load_obj.argtypes = [C.c_char_p, C.c_int]
        L.pto_scene_new.restype = C.c_void_p
        L.pto_scene_new.argtypes = [C.c_void_p, C.c_int]
        L.pto_scene_free.argtypes = [C.c_void_p]
        for f in ("pto_scene_prims", "pto_scene_nodes", "pto_scene_indices"):
            getattr(L, f).restype = C.c_void_p
            getattr(L, f).argtypes = [C.c_void_p]
        L.pto_scene_num_prims.argtypes = [C.c_void_p]
        L.pto_scene_num_nodes.argtypes = [C.c_void_p]
        L.pto_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.pto_primary_ids.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pto_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.pto_render_moments.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.pto_render_init.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.pto_xorwow_init.argtypes = [C.c_uint64, C.c_uint64, C.c_void_p]
        L.pto_camera_lookat.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.pto_tonemap.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.pto_occluded.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.pto_visibility.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.pto_scene_set_grids.argtypes = [C.c_void_p, C.c_void_p, C.c_float]
        L.pto_render_guided.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.pto_scene_get_cdf.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        self.L = L

    # scenes ----------------------------------------------------------------------------------
    def load_obj(self, path, convert_quads=False):
        s = self.L.pto_load_obj(os.fspath(path).encode(), int(convert_quads))
        if not s:
            raise IOError(f"oracle could not load {path}")
        return C.c_void_p(s)

    def scene_from_prims(self, prims):
        a = np.zeros(len(prims), OPRIM_DTYPE)
        for k in ("type", "v", "bsdf", "Le"):
            a[k] = prims[k]
        a["n"] = prims["normal"] if "normal" in prims.dtype.names else prims["n"]
        s = C.c_void_p(self.L.pto_scene_new(_p(a), len(a)))
        if "material_type" in prims.dtype.names and (prims["material_type"] == 1).any():      # extension: mirrors (no reference counterpart)
            flags = np.ascontiguousarray(prims["material_type"] == 1, np.uint8)
            self.L.pto_scene_set_mirrors.argtypes = [C.c_void_p, C.c_void_p]
            self.L.pto_scene_set_mirrors(s, _p(flags))
        return s

    def free(self, s):
        self.L.pto_scene_free(s)

    def prims(self, s):
        n = self.L.pto_scene_num_prims(s)
        buf = (C.c_char * (n * OPRIM_DTYPE.itemsize)).from_address(self.L.pto_scene_prims(s))
        return np.frombuffer(buf, OPRIM_DTYPE).copy()

    def nodes(self, s):
        n = self.L.pto_scene_num_nodes(s)
        buf = (C.c_char * (n * ONODE_DTYPE.itemsize)).from_address(self.L.pto_scene_nodes(s))
        return np.frombuffer(buf, ONODE_DTYPE).copy()

    # camera ----------------------------------------------------------------------------------
    def camera(self, w, h):
        cam = Cam()
        self.L.pto_camera_default(w, h, C.byref(cam))
        return cam

    def camera_lookat(self, eye, at, up, vfov, w, h, orbit=False):
        cam = Cam()
        f3 = C.c_float * 3
        self.L.pto_camera_lookat(f3(*eye), f3(*at), f3(*up), C.c_float(vfov), w, h, int(orbit), C.byref(cam))
        return cam

    # rng -------------------------------------------------------------------------------------
    def xorwow_state(self, seed, subsequence):
        st = XW()
        self.L.pto_xorwow_init(seed, subsequence, C.byref(st))
        return list(st.v) + [st.d]

    def xorwow_uniforms(self, seed, subsequence, n):
        st = XW()
        self.L.pto_xorwow_init(seed, subsequence, C.byref(st))
        return [self.L.pto_uniform(C.byref(st)) for _ in range(n)]

    # queries ---------------------------------------------------------------------------------
    def intersect(self, s, rays, t_min=1e-4, mode=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        ids = np.empty(len(rays), np.int32)
        ts = np.empty(len(rays), np.float32)
        t = C.c_float()
        for i, r in enumerate(rays):
            o = np.ascontiguousarray(r[:3]); d = np.ascontiguousarray(r[3:])
            ids[i] = self.L.pto_intersect(s, _p(o), _p(d), C.c_float(t_min), C.c_float(3.4028234663852886e38), mode, C.byref(t), None)
            ts[i] = t.value
        return ids, ts

    def occluded(self, s, segs, mode=0):
        """Scene::occluded(from, to) for (n, 6) segments; mode 0 = through the BVH2 (as the reference), 1 = linear scan."""
        segs = np.ascontiguousarray(segs, np.float32).reshape(-1, 6)
        out = np.empty(len(segs), np.uint8)
        for i, sg in enumerate(segs):
            a = np.ascontiguousarray(sg[:3]); b = np.ascontiguousarray(sg[3:])
            out[i] = self.L.pto_occluded(s, _p(a), _p(b), mode)
        return out

    def visibility(self, s, w, h, spp, radiosity=None, cam=None, rng=None):
        """render_radiosity: returns (rgb8[h,w,3], rng) -- rng advanced like the kernel does."""
        cam = cam or self.camera(w, h)
        if rng is None:
            rng = self.render_init(w, h)
        rgb = np.zeros((h, w, 3), np.uint8)
        rad = None if radiosity is None else np.ascontiguousarray(radiosity, np.float32)
        self.L.pto_visibility(s, C.byref(cam), rng, spp, None if rad is None else _p(rad), _p(rgb))
        return rgb, rng

    def primary_ids(self, s, w, h, cam=None):
        cam = cam or self.camera(w, h)
        ids = np.empty((h, w), np.int32)
        ts = np.empty((h, w), np.float32)
        self.L.pto_primary_ids(s, C.byref(cam), _p(ids), _p(ts))
        return ids, ts

    def render(self, s, w, h, spp, max_depth, cam=None, rng=None, threads=0):
        """Returns (linear[h,w,3], rgb8[h,w,3], stats[4]).  Pass rng (from render_init) to continue streams."""
        cam = cam or self.camera(w, h)
        if rng is None:
            rng = self.render_init(w, h, threads)
        lin = np.zeros((h, w, 3), np.float32)
        rgb = np.zeros((h, w, 3), np.uint8)
        stats = (C.c_uint64 * 4)()
        self.L.pto_render(s, C.byref(cam), rng, spp, max_depth, _p(lin), _p(rgb), stats, threads)
        return lin, rgb, list(stats)

    def set_grids(self, s, grids_rgb, mis_bsdf_fraction=0.5):
        """grids_rgb: (n_prims, 256, 3) radiosity grids or None (precomputeCDFs, application_state.h:492-585)."""
        g = None if grids_rgb is None else np.ascontiguousarray(grids_rgb, np.float32)
        self.L.pto_scene_set_grids(s, None if g is None else _p(g), C.c_float(mis_bsdf_fraction))

    def cdf(self, s, prim):
        out = np.empty(530, np.float32)
        self.L.pto_scene_get_cdf(s, int(prim), _p(out))
        return out

    def render_guided(self, s, w, h, spp, max_depth, mode, cam=None, rng=None, threads=0):
        """`render` with SamplingMode `mode` (0 BSDF, 3 MIS, 1/2/4 pure grid).  Returns (linear, stats, rng)."""
        cam = cam or self.camera(w, h)
        if rng is None:
            rng = self.render_init(w, h, threads)
        lin = np.zeros((h, w, 3), np.float32)
        stats = (C.c_uint64 * 4)()
        self.L.pto_render_guided(s, C.byref(cam), rng, spp, max_depth, int(mode), _p(lin), stats, threads)
        return lin, list(stats), rng

    def render_init(self, w, h, threads=0):
        rng = (XW * (w * h))()
        self.L.pto_render_init(w, h, rng, threads)
        return rng

    def render_moments(self, s, w, h, spp, max_depth, cam=None, rng=None, threads=0):
        cam = cam or self.camera(w, h)
        if rng is None:
            rng = self.render_init(w, h, threads)
        sm = np.zeros((h, w, 3), np.float64)
        sq = np.zeros((h, w, 3), np.float64)
        self.L.pto_render_moments(s, C.byref(cam), rng, spp, max_depth, _p(sm), _p(sq), threads)
        return sm, sq

    def tonemap(self, linear):
        linear = np.ascontiguousarray(linear, np.float32)
        out = np.empty(linear.shape, np.uint8)
        self.L.pto_tonemap(_p(linear), linear.size // 3, _p(out))
        return out
