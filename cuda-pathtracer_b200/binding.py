"""ctypes binding over the C ABI of include/ptx.h (libptx.so).

Host-side mirror of the reference's scene / camera / integrator interface for test drivers and
bench.py.  It only marshals arguments: every pixel and every hit comes from the CUDA kernels in
libptx.so; if the library is missing or no GPU is usable, calls raise PtxError (no fallback).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PTX_LIB") or os.path.join(HERE, "libptx.so")      # PTX_LIB: experimental builds (tools/)

RNG_XORWOW_COMPAT, RNG_PHILOX = 0, 1
SHARD_NONE, SHARD_TILES, SHARD_SAMPLES = 0, 1, 2
NOALLOC_MAX_RAYS = 65536
RENDER_STAGE_TIMING, RENDER_COUNT_WORK, RENDER_ASYNC, RENDER_SORT_RAYS, RENDER_SORT_MATERIAL, RENDER_NO_COST_ORDER, RENDER_NO_BOUNDS_CULL = 1, 2, 4, 8, 16, 32, 64
PIPELINE_DEFAULT, PIPELINE_FUSED, PIPELINE_LOOP, PIPELINE_PERSIST, PIPELINE_WAVEFRONT, PIPELINE_CTA, PIPELINE_FLAT = 0, 1, 2, 3, 4, 5, 6
SAMPLING_BSDF, SAMPLING_FORMFACTOR, SAMPLING_RADIOSITY, SAMPLING_MIS, SAMPLING_TOPK = 0, 1, 2, 3, 4


class PtxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"ptx error {code}: {msg}")
        self.code = code


class Primitive(C.Structure):
    _fields_ = [("type", C.c_int32), ("v", (C.c_float * 3) * 4), ("normal", C.c_float * 3), ("bsdf", C.c_float * 3),
                ("Le", C.c_float * 3), ("material_type", C.c_int32)]


PRIM_DTYPE = np.dtype([("type", "<i4"), ("v", "<f4", (4, 3)), ("normal", "<f4", 3), ("bsdf", "<f4", 3), ("Le", "<f4", 3),
                       ("material_type", "<i4")])


class Material(C.Structure):
    _fields_ = [("bsdf", C.c_float * 3), ("Le", C.c_float * 3), ("type", C.c_int32), ("reserved", C.c_int32)]


class LoadOpts(C.Structure):
    _fields_ = [("convert_quads", C.c_int32), ("subdivision", C.c_int32), ("triangle_cap", C.c_int64),
                ("mirror_from_mtl", C.c_int32), ("host_preprocess", C.c_int32), ("keep_instances", C.c_int32), ("reserved", C.c_int32)]


class BuildOpts(C.Structure):
    _fields_ = [("builder", C.c_int32), ("sah_bins", C.c_int32), ("max_leaf", C.c_int32), ("reserved", C.c_int32)]


class SceneStats(C.Structure):
    _fields_ = [("n_primitives", C.c_int64), ("n_triangles", C.c_int64), ("n_nodes", C.c_int64), ("n_materials", C.c_int64),
                ("node_bytes", C.c_int64), ("triangle_bytes", C.c_int64), ("shading_bytes", C.c_int64),
                ("bounds_min", C.c_float * 3), ("bounds_max", C.c_float * 3), ("sah_cost", C.c_float), ("build_ms", C.c_float),
                ("builder", C.c_int32), ("max_depth", C.c_int32), ("preprocess_ms", C.c_float), ("preprocess_on_device", C.c_int32),
                ("n_instances", C.c_int32), ("reserved2", C.c_int32), ("stored_triangles", C.c_int64), ("build_stage_ms", C.c_float * 4),
                ("flat_units", C.c_int32), ("reserved", C.c_int32)]


class Camera(C.Structure):
    _fields_ = [("radius", C.c_float), ("yaw", C.c_float), ("pitch", C.c_float), ("vup", C.c_float * 3), ("lookat", C.c_float * 3),
                ("origin", C.c_float * 3), ("lower_left_corner", C.c_float * 3), ("horizontal", C.c_float * 3),
                ("vertical", C.c_float * 3), ("vfov", C.c_float), ("aspect", C.c_float), ("image_width", C.c_int32),
                ("image_height", C.c_int32)]


class RenderOpts(C.Structure):
    _fields_ = [("spp", C.c_int32), ("max_depth", C.c_int32), ("rng_mode", C.c_int32), ("sampling_mode", C.c_int32),
                ("accumulate", C.c_int32), ("shard_mode", C.c_int32), ("shard_rank", C.c_int32), ("shard_world", C.c_int32),
                ("rr_start_depth", C.c_int32), ("rr_cap", C.c_float), ("pipeline", C.c_int32),
                ("flags", C.c_int32), ("loop_iters", C.c_int32), ("ctas_per_sm", C.c_int32), ("chunks", C.c_int32), ("shade_batch", C.c_int32)]


class RenderStats(C.Structure):
    _fields_ = [("segments", C.c_uint64), ("paths", C.c_uint64), ("iterations", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("compactions", C.c_uint64), ("render_ms", C.c_float), ("extend_ms", C.c_float), ("shade_ms", C.c_float),
                ("other_ms", C.c_float), ("node_visits", C.c_uint64), ("tri_tests", C.c_uint64), ("slot_launches", C.c_uint64), ("culled_paths", C.c_uint64)]


EXPORTS = [
    "ptx_ctx_create", "ptx_ctx_destroy", "ptx_last_error", "ptx_sync", "ptx_ctx_device_info", "ptx_ctx_flush_l2", "ptx_version",
    "ptx_timer_start", "ptx_timer_stop", "ptx_ctx_stream",
    "ptx_scene_load", "ptx_parse_scene_file", "ptx_free_primitives", "ptx_scene_from_primitives", "ptx_scene_from_triangles",
    "ptx_scene_build_accel", "ptx_scene_stats_get", "ptx_scene_get_primitives", "ptx_scene_destroy",
    "ptx_scene_set_radiosity_grids", "ptx_scene_get_guide_records",
    "ptx_camera_set", "ptx_camera_update", "ptx_camera_orbit", "ptx_camera_default",
    "ptx_film_create", "ptx_film_reset", "ptx_film_clear", "ptx_film_resolve_rgb8", "ptx_film_read_linear", "ptx_film_write", "ptx_write_png_rgb8",
    "ptx_film_device_buffer", "ptx_film_set_sample_count", "ptx_film_destroy", "ptx_film_save", "ptx_film_load",
    "ptx_shard_pixel_count", "ptx_shard_owned_pixels", "ptx_shard_sample_range",
    "ptx_render", "ptx_render_stats_get", "ptx_comm_unique_id", "ptx_comm_create", "ptx_comm_destroy", "ptx_film_combine",
    "ptx_trace_primary_ids", "ptx_trace_rays", "ptx_trace_rays_noalloc", "ptx_trace_occluded", "ptx_render_visibility", "ptx_film_rng_state",
]

_lib = None


def lib():
    """Load libptx.so (raises if it has not been built: there is no Python/CPU stand-in)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PtxError(-3, f"{LIB_PATH} not built; run __graft_entry__.build()")
        L = C.CDLL(LIB_PATH)
        L.ptx_last_error.restype = C.c_char_p
        L.ptx_last_error.argtypes = [C.c_void_p]
        L.ptx_version.restype = C.c_char_p
        for name in ("ptx_ctx_destroy", "ptx_scene_destroy", "ptx_film_destroy", "ptx_free_primitives"):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [C.c_void_p]
        L.ptx_film_set_sample_count.argtypes = [C.c_void_p, C.c_float]
        L.ptx_ctx_flush_l2.argtypes = [C.c_void_p]
        L.ptx_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_void_p, C.c_void_p]
        L.ptx_comm_unique_id.argtypes = [C.c_void_p, C.c_size_t]
        L.ptx_comm_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.ptx_comm_destroy.argtypes = [C.c_void_p]
        L.ptx_comm_destroy.restype = None
        L.ptx_film_combine.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.ptx_trace_rays_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_void_p, C.c_void_p]
        L.ptx_trace_occluded.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ptx_render_visibility.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.ptx_film_reset.argtypes = [C.c_void_p, C.c_uint64]
        L.ptx_film_rng_state.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
        L.ptx_scene_from_primitives.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.ptx_scene_from_triangles.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t,
                                               C.c_void_p, C.c_void_p]
        L.ptx_scene_get_primitives.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.ptx_scene_set_radiosity_grids.argtypes = [C.c_void_p, C.c_void_p, C.c_float]
        L.ptx_scene_get_guide_records.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
        L.ptx_scene_load.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_void_p]
        L.ptx_parse_scene_file.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ptx_film_write.argtypes = [C.c_void_p, C.c_char_p]
        L.ptx_film_save.argtypes = [C.c_void_p, C.c_char_p]
        L.ptx_film_load.argtypes = [C.c_void_p, C.c_char_p]
        L.ptx_write_png_rgb8.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.ptx_film_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.ptx_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ptx_trace_primary_ids.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.ptx_film_resolve_rgb8.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.ptx_film_read_linear.argtypes = [C.c_void_p, C.c_void_p]
        L.ptx_film_device_buffer.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ptx_camera_set.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float]
        L.ptx_shard_pixel_count.restype = C.c_int64
        L.ptx_shard_pixel_count.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.ptx_shard_owned_pixels.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64]
        L.ptx_shard_sample_range.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _check(rc, ctx=None):
    if rc < 0:
        msg = lib().ptx_last_error(ctx)
        raise PtxError(rc, msg.decode() if msg else "")
    return rc


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def parse_scene_file(path, convert_quads=False, subdivision=0, mirror_from_mtl=False, triangle_cap=0):
    """Host-only parse (loadOBJ / loadPBRT semantics): structured numpy array in the reference's primitive numbering."""
    opts = LoadOpts(int(convert_quads), int(subdivision), int(triangle_cap), int(mirror_from_mtl), 0)
    p = C.c_void_p()
    n = C.c_int64()
    _check(lib().ptx_parse_scene_file(os.fspath(path).encode(), C.byref(opts), C.byref(p), C.byref(n)))
    try:
        buf = (C.c_char * (n.value * C.sizeof(Primitive))).from_address(p.value)
        return np.frombuffer(buf, dtype=PRIM_DTYPE).copy()
    finally:
        lib().ptx_free_primitives(p)


def write_png_rgb8(path, rgb, flip_y=False):
    """Host-only PNG writer (ptx_write_png_rgb8): rgb = (h, w, 3) uint8."""
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    h, w = rgb.shape[:2]
    _check(lib().ptx_write_png_rgb8(os.fspath(path).encode(), int(w), int(h), _ptr(rgb), int(flip_y)))


def shard_owned_pixels(width, height, shard_mode, rank, world):
    """Pixel indices (y*w + x) rank `rank` of `world` renders (host arithmetic only)."""
    n = lib().ptx_shard_pixel_count(int(width), int(height), int(shard_mode), int(rank), int(world))
    _check(int(n))
    out = np.empty(int(n), np.uint32)
    _check(lib().ptx_shard_owned_pixels(int(width), int(height), int(shard_mode), int(rank), int(world), _ptr(out), int(n)))
    return out


def shard_sample_range(spp, shard_mode, rank, world):
    first, count = C.c_int32(), C.c_int32()
    _check(lib().ptx_shard_sample_range(int(spp), int(shard_mode), int(rank), int(world), C.byref(first), C.byref(count)))
    return first.value, count.value


def default_camera(width, height):
    cam = Camera()
    _check(lib().ptx_camera_default(C.byref(cam), int(width), int(height)))
    return cam


def look_at_camera(eye, at, up, vfov, width, height, orbit=False):
    cam = Camera()
    cam.image_width, cam.image_height = int(width), int(height)
    f3 = C.c_float * 3
    _check(lib().ptx_camera_set(C.byref(cam), f3(*eye), f3(*at), f3(*up), C.c_float(vfov), C.c_float(float(width) / float(height))))
    if orbit:
        _check(lib().ptx_camera_orbit(C.byref(cam)))
    return cam


class Context:
    """One CUDA device + stream (ptx_ctx)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        _check(lib().ptx_ctx_create(int(device), C.byref(self._h)))
        self.device = int(device)

    def close(self):
        if self._h:
            lib().ptx_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def sync(self):
        _check(lib().ptx_sync(self._h), self._h)

    def timer_start(self):
        _check(lib().ptx_timer_start(self._h), self._h)

    def timer_stop(self):
        ms = C.c_float()
        _check(lib().ptx_timer_stop(self._h, C.byref(ms)), self._h)
        return ms.value

    def stream(self):
        p = C.c_void_p()
        _check(lib().ptx_ctx_stream(self._h, C.byref(p)), self._h)
        return p.value or 0

    def device_info(self):
        sm, ma, mi, mem = C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
        _check(lib().ptx_ctx_device_info(self._h, C.byref(sm), C.byref(ma), C.byref(mi), C.byref(mem)), self._h)
        return {"sm_count": sm.value, "cc": (ma.value, mi.value), "total_mem": mem.value}

    # -- scenes -------------------------------------------------------------------------------
    def load_scene(self, path, convert_quads=False, subdivision=0, mirror_from_mtl=False, host_preprocess=False, keep_instances=False):
        opts = LoadOpts(int(convert_quads), int(subdivision), 0, int(mirror_from_mtl), int(host_preprocess), int(keep_instances), 0)
        h = C.c_void_p()
        _check(lib().ptx_scene_load(self._h, os.fspath(path).encode(), C.byref(opts), C.byref(h)), self._h)
        return Scene(self, h)

    def scene_from_primitives(self, prims, builder=0, max_leaf=0):
        prims = np.ascontiguousarray(prims, dtype=PRIM_DTYPE)
        bo = BuildOpts(int(builder), 0, int(max_leaf), 0)
        h = C.c_void_p()
        _check(lib().ptx_scene_from_primitives(self._h, _ptr(prims), len(prims), C.byref(bo), C.byref(h)), self._h)
        return Scene(self, h)

    def scene_from_triangles(self, verts, mat_id=None, materials=None, normals=None, builder=0, max_leaf=0):
        verts = np.ascontiguousarray(verts, dtype=np.float32).reshape(-1, 9)
        n = len(verts)
        mid = None if mat_id is None else np.ascontiguousarray(mat_id, dtype=np.uint32)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3)
        mats = (Material * max(len(materials or []), 1))()
        for i, m in enumerate(materials or []):
            mats[i].bsdf[:] = m["bsdf"]; mats[i].Le[:] = m.get("Le", (0, 0, 0)); mats[i].type = int(m.get("type", 0))
        bo = BuildOpts(int(builder), 0, int(max_leaf), 0)
        h = C.c_void_p()
        _check(lib().ptx_scene_from_triangles(self._h, _ptr(verts), None if nrm is None else _ptr(nrm), None if mid is None else _ptr(mid),
                                              mats, n, len(materials or []), C.byref(bo), C.byref(h)), self._h)
        return Scene(self, h)

    def create_film(self, width, height):
        h = C.c_void_p()
        _check(lib().ptx_film_create(self._h, int(width), int(height), C.byref(h)), self._h)
        return Film(self, h, int(width), int(height))

    # -- rendering ----------------------------------------------------------------------------
    def render(self, scene, camera, film, spp, max_depth=5, rng_mode=RNG_XORWOW_COMPAT, accumulate=False, shard_mode=SHARD_NONE,
               shard_rank=0, shard_world=1, pipeline=PIPELINE_DEFAULT, stage_timing=False, count_work=False, sampling_mode=0, loop_iters=0, loop_min_blocks=0, chunks=0, shade_batch=0, async_=False, sort_rays=False, sort_material=False, cost_order=True, bounds_cull=True):
        o = RenderOpts()
        o.spp, o.max_depth, o.rng_mode, o.sampling_mode = int(spp), int(max_depth), int(rng_mode), int(sampling_mode)
        o.accumulate, o.shard_mode, o.shard_rank, o.shard_world = int(accumulate), int(shard_mode), int(shard_rank), int(shard_world)
        o.pipeline = int(pipeline)
        o.flags = (RENDER_STAGE_TIMING if stage_timing else 0) | (RENDER_COUNT_WORK if count_work else 0) | (RENDER_ASYNC if async_ else 0) | (RENDER_SORT_RAYS if sort_rays else 0) | (RENDER_SORT_MATERIAL if sort_material else 0) | (0 if cost_order else RENDER_NO_COST_ORDER) | (0 if bounds_cull else RENDER_NO_BOUNDS_CULL)
        o.loop_iters = int(loop_iters)
        o.ctas_per_sm = int(loop_min_blocks)
        o.chunks = int(chunks)
        o.shade_batch = int(shade_batch)
        _check(lib().ptx_render(self._h, scene._h, C.byref(camera), film._h, C.byref(o)), self._h)
        return None if async_ else self.render_stats()       # async: statistics on the next render_stats() / sync()

    def render_stats(self):
        st = RenderStats()
        _check(lib().ptx_render_stats_get(self._h, C.byref(st)), self._h)
        return {k: getattr(st, k) for k, _ in RenderStats._fields_}

    def trace_primary_ids(self, scene, camera, width, height):
        ids = np.empty((height, width), np.int32)
        ts = np.empty((height, width), np.float32)
        _check(lib().ptx_trace_primary_ids(self._h, scene._h, C.byref(camera), int(width), int(height), _ptr(ids), _ptr(ts)), self._h)
        return ids, ts

    def trace_rays(self, scene, rays, t_min=1e-4):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        ids = np.empty(len(rays), np.int32)
        ts = np.empty(len(rays), np.float32)
        _check(lib().ptx_trace_rays(self._h, scene._h, _ptr(rays), len(rays), C.c_float(t_min), _ptr(ids), _ptr(ts)), self._h)
        return ids, ts

    def create_comm(self, unique_id, rank, world):
        """Collective: every rank passes the 128-byte id rank 0 got from comm_unique_id()."""
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
        h = C.c_void_p()
        _check(lib().ptx_comm_create(self._h, buf, int(rank), int(world), C.byref(h)), self._h)
        return Comm(self, h)

    def trace_rays_noalloc(self, scene, rays, t_min=1e-4):
        """ptx_trace_rays_noalloc: the same query without any device allocation (at most NOALLOC_MAX_RAYS rays per call)."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        ids = np.empty(len(rays), np.int32)
        ts = np.empty(len(rays), np.float32)
        _check(lib().ptx_trace_rays_noalloc(self._h, scene._h, _ptr(rays), len(rays), C.c_float(t_min), _ptr(ids), _ptr(ts)), self._h)
        return ids, ts

    def flush_l2(self):
        _check(lib().ptx_ctx_flush_l2(self._h), self._h)

    def trace_occluded(self, scene, segs):
        """Scene::occluded(from, to) for (n, 6) float32 segments -> uint8[n]."""
        segs = np.ascontiguousarray(segs, dtype=np.float32).reshape(-1, 6)
        out = np.empty(len(segs), np.uint8)
        _check(lib().ptx_trace_occluded(self._h, scene._h, _ptr(segs), len(segs), _ptr(out)), self._h)
        return out

    def render_visibility(self, scene, camera, film, spp, radiosity=None):
        """render_radiosity view: RGB8 (h, w, 3), row 0 = bottom; radiosity = (n_primitives, 3) float32 or None."""
        rad = None if radiosity is None else np.ascontiguousarray(radiosity, dtype=np.float32)
        img = np.empty((film.height, film.width, 3), np.uint8)
        _check(lib().ptx_render_visibility(self._h, scene._h, C.byref(camera), film._h, int(spp), None if rad is None else _ptr(rad), _ptr(img)), self._h)
        return img


def comm_unique_id():
    buf = (C.c_uint8 * 128)()
    _check(lib().ptx_comm_unique_id(buf, 128))
    return bytes(buf)


class Comm:
    def __init__(self, ctx, handle):
        self.ctx, self._h = ctx, handle

    def close(self):
        if self._h:
            lib().ptx_comm_destroy(self._h)
            self._h = C.c_void_p()


class Scene:
    def __init__(self, ctx, handle):
        self.ctx, self._h = ctx, handle

    def close(self):
        if self._h:
            lib().ptx_scene_destroy(self._h)
            self._h = C.c_void_p()

    def stats(self):
        st = SceneStats()
        _check(lib().ptx_scene_stats_get(self._h, C.byref(st)))
        d = {k: getattr(st, k) for k, _ in SceneStats._fields_}
        d["bounds_min"], d["bounds_max"] = list(st.bounds_min), list(st.bounds_max)
        d["build_stage_ms"] = list(st.build_stage_ms)
        return d

    def rebuild(self, builder=0, sah_bins=0, max_leaf=0):
        bo = BuildOpts(int(builder), int(sah_bins), int(max_leaf), 0)
        _check(lib().ptx_scene_build_accel(self._h, C.byref(bo)), self.ctx._h)

    def set_radiosity_grids(self, grids_rgb, mis_bsdf_fraction=0.5):
        """grids_rgb: (n_primitives, 256, 3) float32 radiosity grids (precomputeCDFs input) or None to remove them."""
        g = None if grids_rgb is None else np.ascontiguousarray(grids_rgb, np.float32)
        if g is not None and g.size != self.stats()["n_primitives"] * 768:
            raise ValueError("radiosity grids must hold n_primitives * 256 * 3 floats")
        _check(lib().ptx_scene_set_radiosity_grids(self._h, None if g is None else _ptr(g), C.c_float(mis_bsdf_fraction)), self.ctx._h)

    def guide_records(self, first=0, n=None):
        n = self.stats()["n_primitives"] - first if n is None else n
        out = np.empty((n, 272), np.float32)
        _check(lib().ptx_scene_get_guide_records(self._h, int(first), int(n), _ptr(out)), self.ctx._h)
        return out

    def primitives(self):
        n = _check(lib().ptx_scene_get_primitives(self._h, None, 0))
        out = np.zeros(n, dtype=PRIM_DTYPE)
        if n:
            _check(lib().ptx_scene_get_primitives(self._h, _ptr(out), n))
        return out


class Film:
    def __init__(self, ctx, handle, w, h):
        self.ctx, self._h, self.width, self.height = ctx, handle, w, h

    def close(self):
        if self._h:
            lib().ptx_film_destroy(self._h)
            self._h = C.c_void_p()

    def reset(self, seed_base=2023):
        _check(lib().ptx_film_reset(self._h, int(seed_base)), self.ctx._h)

    def clear(self):
        _check(lib().ptx_film_clear(self._h), self.ctx._h)

    def resolve_device(self, dev_ptr=None, flip_y=False):
        """Resolve into a DEVICE buffer (w*h*3 bytes): caller-owned, or the film's own when dev_ptr is None."""
        _check(lib().ptx_film_resolve_rgb8(self._h, C.c_void_p(dev_ptr) if dev_ptr else None, 1, int(flip_y)), self.ctx._h)

    def rgb8(self, flip_y=False, out=None):
        img = out if out is not None else np.empty((self.height, self.width, 3), np.uint8)
        _check(lib().ptx_film_resolve_rgb8(self._h, _ptr(img), 0, int(flip_y)), self.ctx._h)
        return img

    def linear(self):
        img = np.empty((self.height, self.width, 3), np.float32)
        _check(lib().ptx_film_read_linear(self._h, _ptr(img)), self.ctx._h)
        return img

    def write(self, path):
        _check(lib().ptx_film_write(self._h, os.fspath(path).encode()), self.ctx._h)

    def save(self, path):
        """Checkpoint: running sums, RNG streams and sample counters (ptx_film_save)."""
        _check(lib().ptx_film_save(self._h, os.fspath(path).encode()), self.ctx._h)

    def load(self, path):
        _check(lib().ptx_film_load(self._h, os.fspath(path).encode()), self.ctx._h)

    def device_buffer(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(lib().ptx_film_device_buffer(self._h, C.byref(p), C.byref(n)), self.ctx._h)
        return p.value, n.value

    def combine(self, comm, shard_mode):
        """ptx_film_combine: NCCL exchange of the ranks' films on the context's stream (no host synchronisation)."""
        _check(lib().ptx_film_combine(self._h, comm._h, int(shard_mode)), self.ctx._h)

    def set_sample_count(self, spp_total):
        _check(lib().ptx_film_set_sample_count(self._h, float(spp_total)))

    def rng_state(self, first=0, n=None):
        n = self.width * self.height - first if n is None else n
        out = np.empty((n, 6), np.uint32)
        _check(lib().ptx_film_rng_state(self._h, int(first), int(n), _ptr(out)), self.ctx._h)
        return out
