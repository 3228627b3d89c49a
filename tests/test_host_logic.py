"""CPU tests (-m "not gpu") of the product's host logic and of its kernel cores compiled for the host.

tests/host_emu.cpp compiles traverse.h / shade.h / rng.h (the __host__ __device__ bodies the CUDA
kernels call) plus the host builders and loader into tests/_build/libptx_hostemu.so.  These tests
check them against reference-generated golden vectors and the oracle.  No compute call goes through
libptx.so here (it needs a GPU); for libptx.so only loading and the exported symbols are checked.
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, SCENES


class CamView(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("llc", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("width", C.c_int), ("height", C.c_int)]


@pytest.fixture(scope="module")
def emu(built):
    L = C.CDLL(os.path.join(ROOT, "tests", "_build", "libptx_hostemu.so"))
    L.emu_scene_load.restype = C.c_void_p
    L.emu_scene_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int]
    L.emu_scene_from_prims.restype = C.c_void_p
    L.emu_scene_from_prims.argtypes = [C.c_void_p, C.c_int64, C.c_int]
    L.emu_scene_free.argtypes = [C.c_void_p]
    L.emu_scene_counts.restype = C.c_int64
    L.emu_scene_counts.argtypes = [C.c_void_p] * 5
    L.emu_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    L.emu_primary_ids.argtypes = [C.c_void_p] * 4
    L.emu_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    L.emu_parse_obj.restype = C.c_int64
    L.emu_parse_obj.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int64]
    L.emu_xorwow_init.argtypes = [C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]
    L.emu_occluded.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.emu_inst_load.restype = C.c_void_p
    L.emu_inst_load.argtypes = [C.c_char_p, C.c_int]
    L.emu_inst_free.argtypes = [C.c_void_p]
    L.emu_inst_counts.restype = C.c_int64
    L.emu_inst_counts.argtypes = [C.c_void_p] * 5
    L.emu_inst_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    L.emu_inst_occluded.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.emu_xorwow_skip.argtypes = [C.c_void_p, C.c_uint64]
    L.emu_xorwow_steps.argtypes = [C.c_void_p, C.c_uint64]
    L.emu_cull_rect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    L.emu_flat_units.argtypes = [C.c_void_p, C.c_void_p]
    L.emu_trace_flat.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    L.emu_render_flat.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return L


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "reference_vectors.npz"))


def P(a):
    return a.ctypes.data_as(C.c_void_p)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def cam_view(pkg, w, h):
    c = pkg.default_camera(w, h)
    v = CamView()
    v.origin[:] = c.origin; v.llc[:] = c.lower_left_corner; v.horizontal[:] = c.horizontal; v.vertical[:] = c.vertical
    v.width, v.height = w, h
    return v


# ---- C ABI: the library loads and exports everything include/ptx.h declares ----------------------
def test_abi_exports_every_declared_symbol(pkg):
    hdr = open(os.path.join(ROOT, "include", "ptx.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ptx_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    L = pkg.lib()
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    assert set(pkg.EXPORTS) == declared
    assert b"sm_100a" in L.ptx_version()


def test_struct_sizes_match_header(pkg):
    b = pkg.binding if hasattr(pkg, "binding") else pkg
    assert C.sizeof(b.Primitive) == 4 + 48 + 36 + 4 == b.PRIM_DTYPE.itemsize
    assert C.sizeof(b.Camera) == 4 * 3 + 12 * 6 + 8 + 8
    assert C.sizeof(b.RenderOpts) == 64
    assert C.sizeof(b.Material) == 32


# ---- camera (host maths in libptx.so; no GPU needed) ----------------------------------------------
@pytest.mark.parametrize("w,h", [(512, 512), (1920, 1080), (37, 91)])
def test_default_camera_matches_reference_sensor(pkg, oracle, w, h):
    c = pkg.default_camera(w, h)
    o = oracle.camera(w, h)
    for a, b in ((c.origin, o.origin), (c.lower_left_corner, o.llc), (c.horizontal, o.horizontal), (c.vertical, o.vertical)):
        assert np.array_equal(bits(list(a)), bits(list(b)))
    assert abs(c.origin[0] + 3.7283e-07) < 1e-10 and abs(c.origin[2] - 8.52936) < 1e-5


def test_look_at_camera_matches_oracle(pkg, oracle):
    eye, at, up = (1.0, 2.0, 7.5), (0.2, 2.4, -1.0), (0.1, 1.0, 0.0)
    for orbit in (False, True):
        c = pkg.look_at_camera(eye, at, up, 55.0, 200, 120, orbit=orbit)
        o = oracle.camera_lookat(eye, at, up, 55.0, 200, 120, orbit=orbit)
        for a, b in ((c.origin, o.origin), (c.lower_left_corner, o.llc), (c.horizontal, o.horizontal), (c.vertical, o.vertical)):
            assert np.array_equal(bits(list(a)), bits(list(b)))


# ---- loader ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("tag", ["cbox_cq0", "cbox_quads_cq0", "cbox_quads_cq1"])
def test_obj_loader_matches_reference_loadOBJ(pkg, gold, tag):
    name, cq = tag.rsplit("_cq", 1)
    prims = pkg.parse_scene_file(os.path.join(SCENES, name + ".obj"), convert_quads=bool(int(cq)))
    ref = gold[tag + "_prims"]
    assert len(prims) == len(ref)
    for i, p in enumerate(prims):
        nv = 3 if p["type"] == 0 else 4
        assert p["type"] == ref[i, 0:1].view(np.int32)[0]
        assert np.array_equal(bits(p["v"][:nv].ravel()), bits(ref[i, 1:1 + 3 * nv]))
        assert np.array_equal(bits(p["normal"]), bits(ref[i, 13:16]))
        assert np.array_equal(bits(p["bsdf"]), bits(ref[i, 16:19])) and np.array_equal(bits(p["Le"]), bits(ref[i, 19:22]))


def test_obj_loader_edge_cases(pkg, tmp_path):
    (tmp_path / "m.mtl").write_text("newmtl A\nKd 0.1 0.2 0.3\nKe 1 2 3\nnewmtl M\nKd 0.5 0.5 0.5\nKs 0.9 0.8 0.7\nillum 3\n")
    (tmp_path / "s.obj").write_text(
        "mtllib m.mtl\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nv 0 0 1\nvn 0 0 2\n"
        "f 1 2 3\n"                 # triangle, geometric normal, default material (before usemtl)
        "usemtl A\nf 1//1 2//1 3//1 4//1   # quad with vn and trailing comment\n"
        "f 1/7 2/7 3/7\n"           # v/vt form
        "f 1 2\n"                   # 2 vertices: ignored
        "f 1 2 3 4 5\n"             # 5 vertices: ignored
        "f 1 2 9\n"                 # out of range: ignored
        "f -1 -2 -3\n"              # relative indices: not supported by the reference, ignored
        "usemtl Nope\nf 3 2 1\n"    # unknown material -> default
        "usemtl M\nf 1 2 5\n"
        "s off\no thing\n")
    prims = pkg.parse_scene_file(tmp_path / "s.obj")
    assert [int(p["type"]) for p in prims] == [0, 1, 0, 0, 0]
    assert np.allclose(prims[0]["bsdf"], 0.8) and np.allclose(prims[0]["normal"], [0, 0, 1])
    assert np.allclose(prims[1]["bsdf"], [0.1, 0.2, 0.3]) and np.allclose(prims[1]["Le"], [1, 2, 3]) and np.allclose(prims[1]["normal"], [0, 0, 1])
    assert np.allclose(prims[3]["bsdf"], 0.8) and np.allclose(prims[3]["normal"], [0, 0, -1])
    assert prims[4]["material_type"] == 0 and np.allclose(prims[4]["bsdf"], 0.5)      # reference semantics: Ks/illum ignored
    mir = pkg.parse_scene_file(tmp_path / "s.obj", mirror_from_mtl=True)
    assert mir[4]["material_type"] == 1 and np.allclose(mir[4]["bsdf"], [0.9, 0.8, 0.7])
    sub = pkg.parse_scene_file(tmp_path / "s.obj", subdivision=2)
    assert len(sub) == 5 * 16
    with pytest.raises(pkg.PtxError):
        pkg.parse_scene_file(tmp_path / "missing.obj")
    (tmp_path / "empty.obj").write_text("# nothing\n")
    with pytest.raises(pkg.PtxError):
        pkg.parse_scene_file(tmp_path / "empty.obj")
    with pytest.raises(pkg.PtxError):
        pkg.parse_scene_file(tmp_path / "scene.xyz")


# ---- RNG ------------------------------------------------------------------------------------------
def test_xorwow_jump_tables_reproduce_curand_init(emu):
    import json
    tab = np.zeros(64 * 800, np.uint32)
    emu.emu_jump_tables(P(tab))
    for k in json.load(open(os.path.join(GOLDEN, "xorwow_kat.json"))):
        out = np.zeros(6, np.uint32)
        emu.emu_xorwow_init(k["seed"], k["subsequence"], P(tab), P(out))
        assert list(out) == k["v"] + [k["d"]], k


def test_philox_known_answers(emu):
    """Random123 kat_vectors for philox4x32-10."""
    out = np.zeros(4, np.uint32)
    for ctr_key, want in (((0, 0, 0, 0, 0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
                          ((0xffffffff,) * 6, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
                          ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344, 0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))):
        emu.emu_philox(*[C.c_uint32(x) for x in ctr_key], P(out))
        assert tuple(int(x) for x in out) == want


# ---- BVH8 build + traversal core --------------------------------------------------------------------
@pytest.mark.parametrize("tag", ["cbox_cq0", "cbox_quads_cq0", "cbox_quads_cq1"])
@pytest.mark.parametrize("max_leaf", [1, 3])
def test_bvh8_traversal_primary_ids_match_reference(pkg, emu, gold, tag, max_leaf):
    name, cq = tag.rsplit("_cq", 1)
    s = emu.emu_scene_load(os.path.join(SCENES, name + ".obj").encode(), int(cq), 0, max_leaf)
    for w, h in ((64, 64), (160, 90)):
        cv = cam_view(pkg, w, h)
        ids = np.zeros((h, w), np.int32); ts = np.zeros((h, w), np.float32)
        emu.emu_primary_ids(s, C.byref(cv), P(ids), P(ts))
        assert np.array_equal(ids, gold[f"{tag}_ids_{w}x{h}"].astype(np.int32))
        assert np.array_equal(bits(ts), bits(gold[f"{tag}_t_{w}x{h}"]))
    emu.emu_scene_free(s)


@pytest.mark.parametrize("name", ["cbox", "cbox_quads"])
def test_bvh8_random_rays_match_reference_linear_scan(emu, gold, name):
    s = emu.emu_scene_load(os.path.join(SCENES, name + ".obj").encode(), 0, 0, 3)
    rays = np.ascontiguousarray(gold[name + "_rays"])
    ids = np.zeros(len(rays), np.int32); ts = np.zeros(len(rays), np.float32)
    emu.emu_trace(s, P(rays), len(rays), C.c_float(1e-4), P(ids), P(ts), None)
    ref = gold[name + "_rays_lin_ids"].astype(np.int32)
    diff = ids != ref
    assert diff.sum() <= 2                      # only exact ties may differ
    assert np.array_equal(bits(ts[~diff]), bits(gold[name + "_rays_lin_t"][~diff]))
    emu.emu_scene_free(s)


def test_bvh8_on_subdivided_and_random_soup(pkg, emu, oracle):
    """Deeper trees: subdivided Cornell box (8192 tris) and a random triangle soup, BVH8 vs the oracle's linear scan."""
    rng = np.random.default_rng(5)
    prims = pkg.parse_scene_file(os.path.join(SCENES, "cbox.obj"), subdivision=4)
    assert len(prims) == 32 * 256
    soup = np.zeros(3000, prims.dtype)
    c = rng.uniform(-3, 3, (3000, 1, 3)).astype(np.float32)
    soup["v"][:, :3] = c + rng.normal(scale=0.25, size=(3000, 3, 3)).astype(np.float32)
    soup["normal"] = [0, 0, 1]; soup["bsdf"] = 0.5
    for plist in (prims, soup):
        s = emu.emu_scene_from_prims(P(plist), len(plist), 3)
        nt, nn, dp, sah = C.c_int64(), C.c_int64(), C.c_int(), C.c_float()
        assert emu.emu_scene_counts(s, C.byref(nt), C.byref(nn), C.byref(dp), C.byref(sah)) == len(plist)
        assert nt.value == len(plist) and nn.value >= len(plist) / 24 and dp.value >= 3
        osc = oracle.scene_from_prims(plist)
        o = rng.uniform(-2.5, 2.5, (600, 3)); d = rng.normal(size=(600, 3))
        rays = np.concatenate([o, d], 1).astype(np.float32)
        ids = np.zeros(600, np.int32); ts = np.zeros(600, np.float32); ctr = np.zeros(2, np.uint64)
        emu.emu_trace(s, P(rays), 600, C.c_float(1e-4), P(ids), P(ts), P(ctr))
        oids, ots = oracle.intersect(osc, rays, mode=1)
        same = ids == oids
        assert same.mean() > 0.99
        assert np.array_equal(bits(ts[same]), bits(ots[same]))
        assert ctr[1] / 600 < 0.05 * len(plist)        # the tree actually culls: far fewer tests than a linear scan
        emu.emu_scene_free(s); oracle.free(osc)


def test_empty_and_tiny_scenes(pkg, emu):
    one = np.zeros(1, pkg.PRIM_DTYPE)
    one["v"][0, :3] = [[0, 0, 0], [1, 0, 0], [0, 1, 0]]; one["normal"] = [0, 0, 1]
    rays = np.array([[0.2, 0.2, 1, 0, 0, -1], [2, 2, 1, 0, 0, -1]], np.float32)
    for plist, want in ((one, [0, -1]), (one[:0], [-1, -1])):
        s = emu.emu_scene_from_prims(P(plist) if len(plist) else None, len(plist), 3)
        ids = np.zeros(2, np.int32); ts = np.zeros(2, np.float32)
        emu.emu_trace(s, P(rays), 2, C.c_float(1e-4), P(ids), P(ts), None)
        assert list(ids) == want
        emu.emu_scene_free(s)


# ---- flat unit list for tiny scenes (flat.h) -------------------------------------------------------
@pytest.mark.parametrize("name,cq", [("cbox_quads", 0), ("cbox", 0), ("cbox_quads", 1)])
def test_flat_units_same_hits_as_bvh8(pkg, emu, gold, name, cq):
    """flat_closest (plane + in-plane rectangle filter, one exact test) against the BVH8 traversal and the reference's
    linear scan: same primitive and bit-identical distance for random rays; a whole path-traced frame identical to the bit
    (every path takes the same decisions: equal segment counts), and the every-candidate fallback stays rare."""
    s = emu.emu_scene_load(os.path.join(SCENES, name + ".obj").encode(), cq, 0, 3)
    units = np.zeros((20, 16), np.float32)
    nu = emu.emu_flat_units(s, P(units))
    assert nu == 17                 # 15 planar quads (or coplanar triangle pairs) + the skewed left wall as two triangles
    n = units[:nu, 0:3]
    assert np.allclose(np.linalg.norm(n, axis=1), 1, atol=1e-6)
    assert np.allclose(np.einsum("ij,ij->i", n, units[:nu, 8:11]), 0, atol=1e-6)        # a1 lies in the plane
    assert np.allclose(np.einsum("ij,ij->i", n, units[:nu, 4:7]) - units[:nu, 3], 0, atol=1e-5)   # so does the centre
    if cq == 0:
        rays = np.ascontiguousarray(gold[name + "_rays"])
        ids = np.zeros(len(rays), np.int32); ts = np.zeros(len(rays), np.float32); c3 = np.zeros(3, np.uint64)
        emu.emu_trace_flat(s, P(rays), len(rays), C.c_float(1e-4), C.c_float(0), P(ids), P(ts), P(c3))
        ref = gold[name + "_rays_lin_ids"].astype(np.int32)
        diff = ids != ref
        assert diff.sum() <= 2                      # only exact ties may differ
        assert np.array_equal(bits(ts[~diff]), bits(gold[name + "_rays_lin_t"][~diff]))
        assert c3[0] == len(rays)
    w, h, spp, depth = 96, 54, 6, 16
    v = cam_view(pkg, w, h)
    a = np.zeros((h, w, 3), np.float32); b = np.zeros_like(a)
    sa, sb = C.c_uint64(), C.c_uint64(); c3 = np.zeros(3, np.uint64)
    emu.emu_render(s, C.byref(v), spp, depth, 2023, P(a), None, C.byref(sa))
    emu.emu_render_flat(s, C.byref(v), spp, depth, 2023, P(b), None, C.byref(sb), P(c3))
    assert sa.value == sb.value == c3[0]
    assert np.array_equal(bits(a), bits(b))
    assert c3[1] / c3[0] < 2e-3                      # rays that needed the every-candidate path
    assert c3[2] / c3[0] < 0.7                       # exact unit tests per ray (a third of the camera rays miss everything)
    emu.emu_scene_free(s)


def test_flat_and_bvh8_agree_on_adversarial_tiny_scenes(pkg, emu, oracle):
    """Random scenes of 1..16 quads / triangles (slivers, three sizes, far from the origin) and rays aimed at edges, vertices, started
    on surfaces, axis-parallel, grazing (tests/fuzz_scenes.py).  Both culls are conservative: the flat list's rectangle filter and the
    BVH8 box test over boxes padded like the reference's (bvh_common.h box_pad; rendering/bvh.h:110-150) must never drop a hit that
    the exact triangle test accepts -- so the two intersectors agree bit for bit, and both equal the reference's LINEAR scan.
    (Before the padding the BVH8 path lost 28 of 867 000 such rays -- all aimed at a vertex on a box face -- 23 of which the
    reference's own padded BVH2 keeps.)"""
    import fuzz_scenes
    rng = np.random.default_rng(20250)
    total = lin_checked = 0
    for it in range(60):
        prims, offset, size = fuzz_scenes.tiny_scene(rng, pkg.PRIM_DTYPE)
        s = emu.emu_scene_from_prims(P(prims), len(prims), 3)
        if emu.emu_flat_units(s, None) == 0:
            emu.emu_scene_free(s); continue
        m = 2000
        rays = fuzz_scenes.adversarial_rays(rng, prims, offset, size, m)
        ia = np.zeros(m, np.int32); ta = np.zeros(m, np.float32); ib = np.zeros(m, np.int32); tb = np.zeros(m, np.float32)
        emu.emu_trace(s, P(rays), m, C.c_float(1e-4), P(ia), P(ta), None)
        emu.emu_trace_flat(s, P(rays), m, C.c_float(1e-4), C.c_float(0), P(ib), P(tb), None)
        same_t = bits(ta) == bits(tb)
        assert same_t.all(), (it, rays[~same_t][0], ta[~same_t][0], tb[~same_t][0])
        tie = ia != ib                                  # two primitives at exactly the same distance: either may win
        assert tie.sum() <= 2
        if it % 6 == 0:                                 # the reference's linear scan (python loop per ray: a subset)
            osc = oracle.scene_from_prims(prims)
            d = rays[:300, 3:].astype(np.float32)
            il, tl = oracle.intersect(osc, rays[:300], mode=1)
            ok = (il == ib[:300]) & (bits(np.where(il >= 0, tl, -1.0).astype(np.float32)) == bits(tb[:300]))
            assert (~ok).sum() <= 1, (it, rays[:300][~ok], il[~ok], ib[:300][~ok], tl[~ok], tb[:300][~ok])
            lin_checked += 300
            oracle.free(osc)
        total += m
        emu.emu_scene_free(s)
    assert total >= 100000 and lin_checked >= 2400


def test_bvh8_adversarial_rays_on_clustered_soups(pkg, emu, oracle):
    """Deeper trees: 50..600 primitives in clusters plus an axis-aligned grid of quads (shared edges lying ON box faces), rays aimed
    at edges and vertices.  The padded BVH8 boxes must keep every hit of the reference's linear scan, bit for bit."""
    import fuzz_scenes
    rng = np.random.default_rng(99)
    for it in range(4):
        parts = []
        for _ in range(int(rng.integers(8, 50))):
            p, off, size = fuzz_scenes.tiny_scene(rng, pkg.PRIM_DTYPE)
            p["v"] += (rng.normal(size=3) * 5).astype(np.float32) - off
            parts.append(p)
        g = int(rng.integers(2, 8))
        q = np.zeros(g * g, pkg.PRIM_DTYPE); q["type"] = 1; q["bsdf"] = 0.5; q["normal"] = [0, 1, 0]
        for i in range(g):
            for j in range(g):
                q["v"][i * g + j] = [[i, 0, j], [i + 1, 0, j], [i + 1, 0, j + 1], [i, 0, j + 1]]
        prims = np.concatenate(parts + [q])
        s = emu.emu_scene_from_prims(P(prims), len(prims), 1 + it % 3)
        m = 700
        rays = fuzz_scenes.adversarial_rays(rng, prims, np.zeros(3, np.float32), 3.0, m)
        ia = np.zeros(m, np.int32); ta = np.zeros(m, np.float32)
        emu.emu_trace(s, P(rays), m, C.c_float(1e-4), P(ia), P(ta), None)
        osc = oracle.scene_from_prims(prims)
        il, tl = oracle.intersect(osc, rays, mode=1)
        tl = np.where(il >= 0, tl, -1.0).astype(np.float32)
        assert np.array_equal(bits(ta), bits(tl)), (it, np.nonzero(bits(ta) != bits(tl))[0][:5])
        assert (ia != il).mean() < 0.03                  # same distance to the bit, another primitive: shared edges of the grid
        oracle.free(osc); emu.emu_scene_free(s)


def test_tree_scene_bvh8_vs_reference_goldens(pkg, emu, tmp_path):
    """A 2 566-primitive scene (tests/golden/make_golden_tree.py): the product's .obj loader equals the reference's dump bit for bit on
    a generated file; the BVH8 traversal (host build of the device code) returns the reference's LINEAR-scan hit for every one of
    6 000 adversarial rays (distances bit for bit; ids but for equal-distance ties) -- i.e. also on the 11 rays where the
    reference's own BVH2 disagrees with its linear scan; primary ids equal its BVH's; and the integrator core gives the reference's
    fixed-seed image bit for bit."""
    from cuda_pathtracer_b200 import scenegen
    g = np.load(os.path.join(GOLDEN, "tree_vectors.npz"))
    path = os.path.join(tmp_path, "tree.obj")
    scenegen.write_tree_obj(path, pkg.parse_scene_file(os.path.join(SCENES, "cbox_quads.obj")), level=int(g["level"]))
    prims = pkg.parse_scene_file(path)
    _assert_prims_equal_reference_dump(prims, g["prims"])
    for max_leaf in (1, 3):
        s = emu.emu_scene_from_prims(P(prims), len(prims), max_leaf)
        assert emu.emu_flat_units(s, None) == 0
        rays = np.ascontiguousarray(g["rays"])
        ids = np.zeros(len(rays), np.int32); ts = np.zeros(len(rays), np.float32)
        emu.emu_trace(s, P(rays), len(rays), C.c_float(1e-4), P(ids), P(ts), None)
        assert np.array_equal(bits(ts), bits(g["rays_lin_t"]))
        assert (ids != g["rays_lin"]).sum() <= 30            # same distance to the bit, the neighbouring triangle across a shared edge
        w = h = 160
        v = cam_view(pkg, w, h)
        pid = np.zeros((h, w), np.int32); pt = np.zeros((h, w), np.float32)
        emu.emu_primary_ids(s, C.byref(v), P(pid), P(pt))
        assert np.array_equal(bits(pt), bits(g["ids_t"])) and (pid != g["ids"]).sum() <= 4
        v = cam_view(pkg, 48, 48)
        lin = np.zeros((48, 48, 3), np.float32); seg = C.c_uint64()
        emu.emu_render(s, C.byref(v), 4, 8, 2023, P(lin), None, C.byref(seg))
        assert np.array_equal(bits(lin), bits(g["render_48x48_4spp_d8"]))
        emu.emu_scene_free(s)


def test_xorwow_skip_matrix_equals_stepping(emu):
    """Whole-pixel culling advances a culled pixel's XORWOW stream by 2 * spp draws with T^n (rng.cpp xorwow_build_skip_matrix): the
    state must be the one n single draws leave, for the n of real frames and for awkward ones."""
    rng = np.random.default_rng(3)
    for n in (0, 1, 2, 3, 7, 128, 129, 2 * 64, 2 * 1024, 2 * 4096, 2 * 65535, 1000003):
        a = rng.integers(0, 2**32, 6, dtype=np.uint64).astype(np.uint32)
        b = a.copy()
        emu.emu_xorwow_skip(P(a), n)
        emu.emu_xorwow_steps(P(b), n)
        assert np.array_equal(a, b), n


def test_cull_rectangle_is_conservative(pkg, emu):
    """cull.h: no camera ray through a pixel OUTSIDE the rectangle may hit the box -- checked with the product's own ray generation
    arithmetic (float32, Sensor::get_ray) at the pixel corners and at random jitter, for random cameras and boxes; a camera inside
    or beside the box (a corner behind it) must turn the culling off."""
    rng = np.random.default_rng(11)
    n_rect = n_off = 0
    for it in range(300):
        lo = rng.normal(size=3) * 3
        hi = lo + rng.uniform(0.05, 4.0, 3)
        eye = rng.normal(size=3) * 8 if it % 4 else lo + (hi - lo) * rng.uniform(-0.3, 1.3, 3)       # every fourth camera in or right beside the box
        at = (lo + hi) / 2 + rng.normal(size=3) * rng.choice([0.2, 3.0])
        w, h = int(rng.integers(16, 200)), int(rng.integers(16, 200))
        cam = pkg.look_at_camera(eye, at, [0, 1, 0], float(rng.uniform(15, 80)), w, h)
        out = np.zeros(4, np.int32)
        bmin, bmax = lo.astype(np.float32), hi.astype(np.float32)
        ok = emu.emu_cull_rect(C.byref(cam), P(bmin), P(bmax), w, h, P(out))
        inside = np.all(eye > lo) and np.all(eye < hi)
        if inside:
            assert not ok
        if not ok:
            n_off += 1
            continue
        n_rect += 1
        x0, y0, x1, y1 = out
        assert 0 <= x0 <= x1 <= w and 0 <= y0 <= y1 <= h
        ys, xs = np.mgrid[0:h, 0:w]
        outside = (xs < x0) | (xs >= x1) | (ys < y0) | (ys >= y1)
        px, py = xs[outside].astype(np.float32), ys[outside].astype(np.float32)
        if len(px) == 0:
            continue
        O = np.float32(cam.origin[:]); llc = np.float32(cam.lower_left_corner[:]); hor = np.float32(cam.horizontal[:]); ver = np.float32(cam.vertical[:])
        for ju, jv in [(0, 0), (1, 0), (0, 1), (1, 1)] + [tuple(rng.uniform(0, 1, 2)) for _ in range(4)]:
            u = ((px + np.float32(ju)) / np.float32(w)).astype(np.float32); v = ((py + np.float32(jv)) / np.float32(h)).astype(np.float32)
            d = (llc[None] + u[:, None] * hor[None] + v[:, None] * ver[None] - O[None]).astype(np.float64)
            with np.errstate(divide="ignore", invalid="ignore"):
                t0 = (bmin.astype(np.float64)[None] - O[None]) / d; t1 = (bmax.astype(np.float64)[None] - O[None]) / d
            tn = np.nanmax(np.minimum(t0, t1), axis=1); tf = np.nanmin(np.maximum(t0, t1), axis=1)
            hit = (tn <= tf) & (tf > 0)
            assert not hit.any(), (it, out, np.nonzero(hit)[0][:4])
    assert n_rect > 100 and n_off > 20, (n_rect, n_off)


def test_flat_units_reject_large_or_degenerate(pkg, emu):
    prims = pkg.parse_scene_file(os.path.join(SCENES, "cbox.obj"), subdivision=1)      # 128 triangles
    s = emu.emu_scene_from_prims(P(prims), len(prims), 3)
    assert emu.emu_flat_units(s, None) == 0
    emu.emu_scene_free(s)
    one = np.zeros(2, pkg.PRIM_DTYPE)
    one["v"][0, :3] = [[0, 0, 0], [1, 0, 0], [0, 1, 0]]; one["v"][1, :3] = [[0, 0, 0], [1, 0, 0], [2, 0, 0]]   # zero area
    s = emu.emu_scene_from_prims(P(one), 2, 3)
    assert emu.emu_flat_units(s, None) == 0
    emu.emu_scene_free(s)
    s = emu.emu_scene_from_prims(P(one[:1]), 1, 3)                                       # a single triangle: one unit, null second record
    assert emu.emu_flat_units(s, None) == 1
    rays = np.array([[0.2, 0.2, 1, 0, 0, -1], [2, 2, 1, 0, 0, -1], [0.2, 0.2, 1, 1, 0, 0]], np.float32)
    ids = np.zeros(3, np.int32); ts = np.zeros(3, np.float32)
    emu.emu_trace_flat(s, P(rays), 3, C.c_float(1e-4), C.c_float(0), P(ids), P(ts), None)
    assert list(ids) == [0, -1, -1] and ts[0] == 1.0
    emu.emu_scene_free(s)


def test_negative_zero_direction(emu):
    """A direction component of -0.0f: round 1's Traversal::begin took 1/d negative (copysign) while visit() ordered the slab
    planes as for a positive component, so such a ray missed everything.  Both intersectors must agree with +0.0f."""
    s = emu.emu_scene_load(os.path.join(SCENES, "cbox_quads.obj").encode(), 0, 0, 3)
    base = [2.54588914, 5.31591082, -4.15390921]
    rays = np.array([base + [-0.0, -0.999709547, -0.0241003428], base + [0.0, -0.999709547, -0.0241003428],
                     [0, 2.5, 3, -0.0, -0.0, -1], [0, 2.5, 3, 0.0, 0.0, -1], [0, 5.0, -2, 0.3, -0.0, -0.0], [0, 5.0, -2, 0.3, 0.0, 0.0]], np.float32)
    for fn in ("bvh", "flat"):
        ids = np.zeros(len(rays), np.int32); ts = np.zeros(len(rays), np.float32)
        if fn == "bvh":
            emu.emu_trace(s, P(rays), len(rays), C.c_float(1e-4), P(ids), P(ts), None)
        else:
            emu.emu_trace_flat(s, P(rays), len(rays), C.c_float(1e-4), C.c_float(0), P(ids), P(ts), None)
        assert ids[0] == ids[1] == 3 and ts[0] == ts[1] > 5
        assert ids[2] == ids[3] >= 0 and ts[2] == ts[3]
        assert ids[4] == ids[5] >= 0 and ts[4] == ts[5]
    emu.emu_scene_free(s)


def test_mirror_extension_core_vs_oracle(pkg, emu, oracle):
    """Mirrors are an extension (the reference integrator is Lambertian-only): the oracle restates the product's own rule, and the
    shading core compiled for the host must agree with it bit for bit on C4's geometry (tall box mirror, short box emissive)."""
    from cuda_pathtracer_b200 import scenegen
    prims = scenegen.c4_prims(pkg.parse_scene_file(os.path.join(SCENES, "cbox_quads.obj")))
    w, h, spp, depth = 64, 36, 4, 8
    osc = oracle.scene_from_prims(prims)
    olin, _, ostats = oracle.render(osc, w, h, spp, depth)
    s = emu.emu_scene_from_prims(P(prims), len(prims), 3)
    v = cam_view(pkg, w, h)
    lin = np.zeros((h, w, 3), np.float32); seg = C.c_uint64()
    emu.emu_render(s, C.byref(v), spp, depth, 2023, P(lin), None, C.byref(seg))
    assert seg.value == ostats[0]
    assert np.array_equal(bits(lin), bits(olin))
    emu.emu_scene_free(s); oracle.free(osc)


# ---- two-level instancing (traverse_inst.h, instancing_host.h) ------------------------------------
def test_two_level_traversal_matches_flattened_scene(pkg, emu, tmp_path):
    """keep_instances on the CPU: the product's instanced loader, instance records (instancing_host.h) and InstTraversal compiled for
    the host, trees from the host builder -- against the FLATTENED load of the same file (what the reference's loadPBRT produces,
    pinned bit for bit by pbrt_vectors.npz) traversed single-level.  Same primitive numbering, distances equal to rounding, and the
    shading normal resolved per instance equals the flattened primitive's stored normal."""
    from cuda_pathtracer_b200 import scenegen
    path = str(tmp_path / "inst.pbrt")
    n = scenegen.write_c5_pbrt(path, instances=20, level=3, n_lights=4)
    flat_prims = pkg.parse_scene_file(path, mirror_from_mtl=True)
    assert len(flat_prims) == n
    sf = emu.emu_scene_from_prims(P(flat_prims), len(flat_prims), 3)
    si = emu.emu_inst_load(path.encode(), 1)
    assert si
    ni, stored, dt, db = C.c_int64(), C.c_int64(), C.c_int(), C.c_int()
    assert emu.emu_inst_counts(si, C.byref(ni), C.byref(stored), C.byref(dt), C.byref(db)) == n
    assert ni.value == 21 and stored.value == 2 * 20 * 4 ** 3 + 10 + 2 * 4
    assert 2 * dt.value + 1 + db.value <= 24                                    # fits the kernels' traversal stack
    rng = np.random.default_rng(9)
    o = rng.uniform(0.5, 11.5, (4000, 3)); d = rng.normal(size=(4000, 3))
    rays = np.concatenate([o, d], 1).astype(np.float32)
    idf = np.zeros(4000, np.int32); tf = np.zeros(4000, np.float32)
    idi = np.zeros(4000, np.int32); ti = np.zeros(4000, np.float32); nrm = np.zeros((4000, 3), np.float32)
    emu.emu_trace(sf, P(rays), 4000, C.c_float(1e-4), P(idf), P(tf), None)
    emu.emu_inst_trace(si, P(rays), 4000, C.c_float(1e-4), P(idi), P(ti), P(nrm))
    assert (idf >= 0).mean() > 0.8                                               # the room is open at the front
    same = idf == idi
    assert same.mean() >= 0.999, float(same.mean())
    close = np.isclose(ti[same], tf[same], rtol=1e-4, atol=1e-4)
    assert close.mean() >= 0.999
    hit = same & (idf >= 0)
    want = flat_prims["normal"][idf[hit]]
    assert np.abs(nrm[hit] - want).max() < 2e-3                                  # unit(L e1 x L e2) against the flattened unit((v1 - v0) x (v2 - v0))
    # any-hit (Scene::occluded) through the two-level structure: the same booleans as the flattened scene, but for segments whose
    # end point lies within rounding of a surface (object-space arithmetic)
    a = rng.uniform(0.5, 11.5, (6000, 3)); b = rng.uniform(0.5, 11.5, (6000, 3))
    hitp = rays[hit][:1500, :3] + rays[hit][:1500, 3:] / np.linalg.norm(rays[hit][:1500, 3:], axis=1, keepdims=True) * tf[hit][:1500, None]
    segs = np.concatenate([np.concatenate([a, b], 1), np.concatenate([rays[hit][:1500, :3], hitp], 1)], 0).astype(np.float32)    # the last 1500 end ON a surface
    of = np.zeros(len(segs), np.uint8); oi = np.zeros(len(segs), np.uint8)
    emu.emu_occluded(sf, P(segs), len(segs), P(of))
    emu.emu_inst_occluded(si, P(segs), len(segs), P(oi))
    assert 0.2 < of[:6000].mean() < 0.98
    assert (of[:6000] != oi[:6000]).mean() < 1e-3
    assert (of[6000:] != oi[6000:]).mean() < 0.2                                  # borderline by construction: rounding decides
    emu.emu_scene_free(sf); emu.emu_inst_free(si)


# ---- any-hit end-point rule (traverse.h trace_any) -------------------------------------------------
@pytest.mark.parametrize("name,cq", [("cbox", 0), ("cbox_quads", 0), ("cbox_quads", 1)])
def test_occluded_endpoint_rule_matches_reference(pkg, emu, oracle, name, cq):
    """Segments ending within a few ulps of a surface (tests/golden/f4_boundary.npz, answers from the reference's own
    Scene::occluded): a triangle accepts t == t_max, the halves of a quad do not (triangle.h:83, quad.h:76).  The oracle
    restatement and the product's trace_any (compiled for the host) must both reproduce every answer."""
    g = np.load(os.path.join(GOLDEN, "f4_boundary.npz"))
    segs = np.ascontiguousarray(g["segs"]); want = g[f"{name}_cq{cq}_occ"]
    assert np.array_equal(want, g[f"{name}_cq{cq}_occ_lin"])
    assert (g["cbox_quads_cq0_occ"] != g["cbox_quads_cq1_occ"]).sum() > 100          # the vectors do contain t == t_max cases
    path = os.path.join(SCENES, name + ".obj")
    osc = oracle.load_obj(path, convert_quads=bool(cq))
    sub = slice(0, len(segs), 5)
    assert np.array_equal(oracle.occluded(osc, segs[sub], mode=0), want[sub])
    oracle.free(osc)
    s = emu.emu_scene_load(path.encode(), cq, 0, 3)
    out = np.zeros(len(segs), np.uint8)
    emu.emu_occluded(s, P(segs), len(segs), P(out))
    assert np.array_equal(out, want), int((out != want).sum())
    emu.emu_scene_free(s)


# ---- integrator core (shade.h) --------------------------------------------------------------------
@pytest.mark.parametrize("tag,w,h,spp,depth", [("cbox_cq0", 48, 48, 8, 8), ("cbox_quads_cq0", 64, 36, 4, 16), ("cbox_quads_cq1", 32, 32, 16, 5)])
def test_integrator_core_bit_exact_vs_reference(pkg, emu, gold, tag, w, h, spp, depth):
    name, cq = tag.rsplit("_cq", 1)
    s = emu.emu_scene_load(os.path.join(SCENES, name + ".obj").encode(), int(cq), 0, 3)
    cv = cam_view(pkg, w, h)
    lin = np.zeros((h, w, 3), np.float32); rgb = np.zeros((h, w, 3), np.uint8); seg = C.c_uint64()
    emu.emu_render(s, C.byref(cv), spp, depth, 2023, P(lin), P(rgb), C.byref(seg))
    key = f"{tag}_{w}x{h}_s{spp}_d{depth}"
    assert np.array_equal(bits(lin), bits(gold[key + "_linear"]))
    assert np.array_equal(rgb, gold[key + "_rgb8"])
    emu.emu_scene_free(s)


@pytest.mark.parametrize("name", ["basic", "materials", "transforms", "plymesh"])
def test_pbrt_loader_matches_reference_loadPBRT(pkg, name):
    """tests/golden/pbrt_vectors.npz holds what the reference's loadPBRT (utils/pbrt_loader.h:178-422 + its vendored
    pbrtParser) returns for tests/pbrt/*.pbrt (tests/golden/make_golden_pbrt.py).  Order, vertices, normals, albedo and
    emission must match bit for bit."""
    gold = np.load(os.path.join(GOLDEN, "pbrt_vectors.npz"))
    pr = pkg.parse_scene_file(os.path.join(ROOT, "tests", "pbrt", name + ".pbrt"))
    assert len(pr) == len(gold[name + "_v"])
    assert (pr["type"] == 0).all() and (gold[name + "_type"] == 0).all()
    assert np.array_equal(pr["v"][:, :3].reshape(len(pr), 9), gold[name + "_v"])
    assert np.array_equal(pr["normal"], gold[name + "_normal"])
    assert np.array_equal(pr["bsdf"], gold[name + "_bsdf"])
    assert np.array_equal(pr["Le"], gold[name + "_Le"])
    assert (pr["material_type"] == 0).all()                      # reference semantics: mirrors are diffuse albedos


def test_pbrt_loader_extension_and_errors(pkg, tmp_path):
    pr = pkg.parse_scene_file(os.path.join(ROOT, "tests", "pbrt", "materials.pbrt"), mirror_from_mtl=True)
    assert list(np.nonzero(pr["material_type"] == 1)[0]) == [4, 5]            # the two `Material "mirror"` triangles
    bad = tmp_path / "bad.pbrt"
    bad.write_text('WorldBegin\nShape "plymesh" "string filename" "x.ply"\nWorldEnd\n')       # missing file
    with pytest.raises(pkg.PtxError):
        pkg.parse_scene_file(str(bad))
    # a PLY file with a quad face: the reference's parser throws "only triangles are supported" and loadPBRT fails
    quad = os.path.join(ROOT, "tests", "pbrt", "quad_ascii.ply")
    bad.write_text(f'WorldBegin\nShape "plymesh" "string filename" "../../../../../../../../{quad}"\n'
                   'Shape "trianglemesh" "point P" [0 0 0 1 0 0 0 1 0] "integer indices" [0 1 2]\nWorldEnd\n')
    with pytest.raises(pkg.PtxError) as e:
        pkg.parse_scene_file(str(bad))
    assert "only triangles are supported" in str(e.value)
    bad.write_text('WorldBegin\nFrobnicate 1 2 3\nWorldEnd\n')
    with pytest.raises(pkg.PtxError):
        pkg.parse_scene_file(str(bad))
    bad.write_text('WorldBegin\nShape "sphere" "float radius" [1]\nWorldEnd\n')      # no triangles: loadPBRT returns false
    with pytest.raises(pkg.PtxError):
        pkg.parse_scene_file(str(bad))
    with pytest.raises(pkg.PtxError):
        pkg.parse_scene_file(str(tmp_path / "missing.pbrt"))


def test_pbrt_indices_above_2_pow_24(pkg, tmp_path):
    """"integer indices" are parsed as integers: float32 is exact only up to 2^24, so a mesh with more than 16.7 M vertices
    used to get its indices silently rounded (2^24 + 1 -> 2^24) and rendered the wrong geometry.  Also: Include cycles fail."""
    nv = 2 ** 24 + 4
    f = tmp_path / "big.pbrt"
    with open(f, "w") as fh:
        fh.write('WorldBegin\nShape "trianglemesh" "point P" [\n')
        chunk = "0 0 0 " * 65536 + "\n"
        for _ in range((nv - 4) // 65536):
            fh.write(chunk)
        fh.write("0 0 0 " * ((nv - 4) % 65536))
        fh.write("1 1 5  2 1 5  2 2 5  1 2 5 ]\n")
        fh.write(f'"integer indices" [ {nv - 4} {nv - 3} {nv - 2}  {nv - 4} {nv - 2} {nv - 1} ]\nWorldEnd\n')
    prims = pkg.parse_scene_file(str(f))
    assert len(prims) == 2
    assert prims["v"][0, :3].tolist() == [[1, 1, 5], [2, 1, 5], [2, 2, 5]]
    assert prims["v"][1, :3].tolist() == [[1, 1, 5], [2, 2, 5], [1, 2, 5]]
    os.remove(f)
    loop = tmp_path / "loop.pbrt"
    loop.write_text('WorldBegin\nInclude "loop.pbrt"\nWorldEnd\n')
    with pytest.raises(pkg.PtxError) as e:
        pkg.parse_scene_file(str(loop))
    assert "Include nested deeper" in str(e.value)


def test_pbrt_triangle_cap_proxy(pkg):
    """ptx_load_opts.triangle_cap reproduces the reference's bounding-box proxy (pbrt_loader.h:205-272): 12 triangles,
    albedo (0.8, 0.2, 0.2), spanning the scene bounds."""
    path = os.path.join(ROOT, "tests", "pbrt", "basic.pbrt")
    full = pkg.parse_scene_file(path)
    proxy = pkg.parse_scene_file(path, triangle_cap=5)
    assert len(proxy) == 12 and np.allclose(proxy["bsdf"], [0.8, 0.2, 0.2])
    fv, pv = full["v"][:, :3].reshape(-1, 3), proxy["v"][:, :3].reshape(-1, 3)
    assert np.array_equal(fv.min(0), pv.min(0)) and np.array_equal(fv.max(0), pv.max(0))
    assert len(pkg.parse_scene_file(path, triangle_cap=14)) == 14


def _decode_png(path):
    """Minimal PNG reader for the test (8-bit RGB, filter 0 only): checks signature, chunk CRCs and the zlib stream."""
    import struct, zlib
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])
        assert crc == zlib.crc32(typ + data) & 0xFFFFFFFF
        chunks.append((typ, data)); pos += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    w, h, depth, ctype, comp, flt, inter = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, ctype, comp, flt, inter) == (8, 2, 0, 0, 0)
    raw = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(h, 1 + 3 * w)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 3)


@pytest.mark.parametrize("w,h", [(1, 1), (7, 5), (301, 233)])
def test_png_writer_roundtrip(pkg, tmp_path, w, h):
    """The film's PNG export ("Save PNG", ui_windows.h:195-210): valid chunks / CRC / Adler, pixels intact, and flip_y
    writing the rows bottom-up like stbi_flip_vertically_on_write(1).  (301 x 233 x 3 spans several stored blocks.)"""
    img = np.random.default_rng(w * h).integers(0, 256, (h, w, 3), dtype=np.uint8)
    pkg.write_png_rgb8(tmp_path / "a.png", img)
    assert np.array_equal(_decode_png(tmp_path / "a.png"), img)
    pkg.write_png_rgb8(tmp_path / "b.png", img, flip_y=True)
    assert np.array_equal(_decode_png(tmp_path / "b.png"), img[::-1])
    with pytest.raises(pkg.PtxError):
        pkg.write_png_rgb8(tmp_path / "nodir" / "c.png", img)


def _assert_prims_equal_reference_dump(prims, ref):
    """prims: structured array (binding.PRIM_DTYPE); ref: (n, 22) float32 rows of the reference harness dump."""
    assert len(prims) == len(ref)
    ty = ref[:, 0].copy().view(np.int32)
    assert np.array_equal(prims["type"], ty)
    v = prims["v"][:, :, :3].reshape(len(prims), 12)
    tri = ty == 0
    assert np.array_equal(bits(v[tri, :9]), bits(ref[tri, 1:10])) and np.array_equal(bits(v[~tri]), bits(ref[~tri, 1:13]))
    assert np.array_equal(bits(prims["normal"]), bits(ref[:, 13:16]))
    assert np.array_equal(bits(prims["bsdf"]), bits(ref[:, 16:19])) and np.array_equal(bits(prims["Le"]), bits(ref[:, 19:22]))


@pytest.mark.parametrize("tag", ["cbox_cq0_sub1", "cbox_cq0_sub3", "cbox_quads_cq0_sub2", "cbox_quads_cq1_sub1", "cbox_quads_cq1_sub2"])
def test_host_subdivision_matches_reference(pkg, tag):
    """Host convert-quads + subdivision (csrc/loader.cpp) against the reference's convertQuadsToTriangles + subdivide_primitives
    (tests/golden/f3_vectors.npz): order, vertices and recomputed normals bit for bit."""
    g = np.load(os.path.join(GOLDEN, "f3_vectors.npz"))
    name, rest = tag.split("_cq")
    cq, sub = rest.split("_sub")
    prims = pkg.parse_scene_file(os.path.join(SCENES, name + ".obj"), convert_quads=bool(int(cq)), subdivision=int(sub))
    _assert_prims_equal_reference_dump(prims, g[tag])


def test_pbrt_loader_generated_instanced_scene(pkg, tmp_path):
    """A generated scene of the C5 shape (scenegen.write_c5_pbrt: object definitions, 27 instances under Translate / Rotate /
    Scale, area lights) must flatten to exactly what the reference's loadPBRT produces: count, SHA-256 of every array and
    every 61st primitive are stored in tests/golden/pbrt_vectors.npz (make_golden_pbrt.py)."""
    import hashlib
    from cuda_pathtracer_b200 import scenegen
    g = np.load(os.path.join(GOLDEN, "pbrt_vectors.npz"))
    path = tmp_path / "c5_small.pbrt"
    n = scenegen.write_c5_pbrt(str(path), instances=27, level=3, n_lights=4)
    pr = pkg.parse_scene_file(path)
    assert len(pr) == n == int(g["c5_small_count"][0])
    arrays = {"v": pr["v"][:, :3].reshape(len(pr), 9), "normal": pr["normal"], "bsdf": pr["bsdf"], "Le": pr["Le"]}
    for k, a in arrays.items():
        a = np.ascontiguousarray(a)
        assert np.array_equal(a[::61], g["c5_small_" + k + "_sampled"]), k
        assert hashlib.sha256(a.tobytes()).digest() == g["c5_small_" + k + "_sha256"].tobytes(), k
    mir = pkg.parse_scene_file(path, mirror_from_mtl=True)
    assert (mir["material_type"] == 1).sum() == 4 * 1280          # instances 0, 7, 14, 21 use the mirror definition


# ---- C++ shim in the reference's vocabulary (include/ptx/compat.hpp) + headless render loop (examples/) -------------
def test_cpp_compat_shim_builds_and_mirrors_sensor(pkg, built):
    """examples/render_loop.cpp is written only against include/ptx/compat.hpp (Sensor, RenderState, SceneState,
    renderFrame with the reference's names).  It must compile and link against libptx.so, its Sensor must reproduce the
    reference camera (sensor.h:38-67) bit for bit, and without a GPU the loop must fail loudly (no CPU rendering path)."""
    import subprocess
    exe = built.build_example()
    r = subprocess.run([exe, "--camera", "1920", "1080"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    vals = np.array([float(x) for x in r.stdout.split()], np.float32)
    c = pkg.default_camera(1920, 1080)
    want = np.array(list(c.origin) + list(c.lower_left_corner) + list(c.horizontal) + list(c.vertical) + [c.radius], np.float32)
    assert np.array_equal(bits(vals), bits(want))          # %.9g round-trips a float32
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:
        r = subprocess.run([exe, os.path.join(SCENES, "cbox.obj"), "32", "32", "1", "1", "-"], capture_output=True, text=True)
        assert r.returncode == 3 and "no CPU path" in r.stderr
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


# ---- grid-guided / MIS sampling core (shade.h guided_direction, build_guide_record), compiled for the CPU ------------
def _guided_meta():
    import json
    return json.load(open(os.path.join(GOLDEN, "guided_vectors.json")))


def _synthetic_grids(n):
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_guided", os.path.join(GOLDEN, "make_golden_guided.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.synthetic_grids(n)


def test_guide_records_match_reference_cdfs(emu, oracle):
    """build_guide_record (the body of the k_build_guide kernel) against the oracle's restatement of Grid::initFromRadiosity
    + buildCDFs, itself pinned through the reference's rendered images: pdf, conditional CDFs, marginal CDF, total weight
    and validity bit for bit, including empty grids, empty rows and empty cells."""
    emu.emu_build_guide.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    s = oracle.load_obj(os.path.join(SCENES, "cbox.obj"), 0)
    grids = _synthetic_grids(32)
    oracle.set_grids(s, grids, 0.5)
    rec = np.zeros((32, 272), np.float32)
    emu.emu_build_guide(P(grids), 32, P(rec))
    for p in range(32):
        c = oracle.cdf(s, p)                           # pdf[256] row_sums[8] marginal[8] row_cdfs[256] total valid
        assert np.array_equal(bits(rec[p, 0:128]), bits(c[0:128]))
        assert np.array_equal(bits(rec[p, 128:256]), bits(c[272:400]))
        assert np.array_equal(bits(rec[p, 256:264]), bits(c[264:272]))
        assert bits(rec[p, 264:265])[0] == bits(c[528:529])[0]
        assert rec[p, 265:266].view(np.uint32)[0] == int(c[529])
    assert {int(v) for v in rec[:, 265].view(np.uint32)} == {0, 1}
    oracle.free(s)


@pytest.mark.parametrize("tag", sorted(_guided_meta()))
def test_guided_core_bit_exact_vs_reference(pkg, emu, tag):
    """The device shading core with grid / MIS sampling, compiled for the host, against the reference's own integrator run
    on the host (tests/golden/guided_vectors.npz): radiance and final RNG states bit for bit."""
    emu.emu_build_guide.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    emu.emu_render_guided.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_float, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p]
    m = _guided_meta()[tag]
    gold = np.load(os.path.join(GOLDEN, "guided_vectors.npz"))
    s = emu.emu_scene_load(os.path.join(SCENES, m["scene"]).encode(), int(m["convert_quads"]), 0, 3)
    grids = _synthetic_grids(m["n_prims"])
    rec = np.zeros((m["n_prims"], 272), np.float32)
    emu.emu_build_guide(P(grids), m["n_prims"], P(rec))
    w, h = m["w"], m["h"]
    cv = cam_view(pkg, w, h)
    lin = np.zeros((h, w, 3), np.float32); st = np.zeros((h * w, 6), np.uint32); seg = C.c_uint64()
    emu.emu_render_guided(s, C.byref(cv), m["spp"], m["depth"], 2023, m["mode"], C.c_float(m["bsdf_fraction"]), P(rec), P(lin), P(st), C.byref(seg))
    assert np.array_equal(st, gold[tag + "_rng"])
    assert np.array_equal(bits(lin), bits(gold[tag + "_linear"]))
    emu.emu_scene_free(s)
