"""CPU tests (-m "not gpu"): pin the plain-C oracle to vectors produced by the reference's own code.

tests/golden/reference_vectors.npz and xorwow_kat.json come from oracle/_ref/ref_host, i.e. the
UNMODIFIED headers of /root/reference compiled for the host (tests/golden/make_golden.py).  The
oracle must reproduce them BIT FOR BIT; that is what "parity pinned" in oracle/pt_oracle.h means.
"""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, SCENES


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "reference_vectors.npz"))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_xorwow_known_answers(oracle):
    kats = json.load(open(os.path.join(GOLDEN, "xorwow_kat.json")))
    assert len(kats) >= 8
    for k in kats:
        assert oracle.xorwow_state(k["seed"], k["subsequence"]) == k["v"] + [k["d"]], k
        got = oracle.xorwow_uniforms(k["seed"], k["subsequence"], len(k["uniform"]))
        assert np.array_equal(bits(got), bits(k["uniform"])), k
    # survey KAT: curand_init(2030, 7, 0)
    u = oracle.xorwow_uniforms(2030, 7, 4)
    assert np.allclose(u, [0.450530142, 0.219993725, 0.328868538, 0.271020651], rtol=0, atol=1e-9)


@pytest.mark.parametrize("tag", ["cbox_cq0", "cbox_quads_cq0", "cbox_quads_cq1"])
def test_loader_and_bvh_builder_match_reference(oracle, gold, tag):
    name, cq = tag.rsplit("_cq", 1)
    s = oracle.load_obj(os.path.join(SCENES, name + ".obj"), int(cq))
    ref = gold[tag + "_prims"]
    prims = oracle.prims(s)
    assert len(prims) == len(ref) == {"cbox_cq0": 32, "cbox_quads_cq0": 16, "cbox_quads_cq1": 32}[tag]
    assert np.array_equal(prims["type"], ref[:, 0].view(np.int32))
    for i, p in enumerate(prims):
        nv = 3 if p["type"] == 0 else 4
        assert np.array_equal(bits(p["v"][:nv].ravel()), bits(ref[i, 1:1 + 3 * nv]))
        assert np.array_equal(bits(p["n"]), bits(ref[i, 13:16]))
        assert np.array_equal(bits(p["bsdf"]), bits(ref[i, 16:19]))
        assert np.array_equal(bits(p["Le"]), bits(ref[i, 19:22]))
    nodes = oracle.nodes(s)
    rn = gold[tag + "_nodes"]
    assert len(nodes) == len(rn)
    assert np.array_equal(bits(nodes["bmin"]), bits(rn[:, 0:3])) and np.array_equal(bits(nodes["bmax"]), bits(rn[:, 3:6]))
    assert np.array_equal(nodes["left"], rn[:, 6].view(np.int32)) and np.array_equal(nodes["count"], rn[:, 8].view(np.int32))
    oracle.free(s)


@pytest.mark.parametrize("tag", ["cbox_cq0", "cbox_quads_cq0", "cbox_quads_cq1"])
@pytest.mark.parametrize("w,h", [(64, 64), (160, 90)])
def test_primary_ids_bit_exact(oracle, gold, tag, w, h):
    name, cq = tag.rsplit("_cq", 1)
    s = oracle.load_obj(os.path.join(SCENES, name + ".obj"), int(cq))
    ids, ts = oracle.primary_ids(s, w, h)
    assert np.array_equal(ids, gold[f"{tag}_ids_{w}x{h}"].astype(np.int32))
    assert np.array_equal(bits(ts), bits(gold[f"{tag}_t_{w}x{h}"]))
    if tag == "cbox_cq0" and w == 64:        # SURVEY.md A.2 KAT
        assert (ids >= 0).sum() == 3080 and ids[ids >= 0].sum() == 28699
    oracle.free(s)


@pytest.mark.parametrize("tag", ["cbox_cq0", "cbox_quads_cq0", "cbox_quads_cq1"])
@pytest.mark.parametrize("w,h,spp,depth", [(48, 48, 8, 8), (64, 36, 4, 16), (32, 32, 16, 5)])
def test_render_bit_exact(oracle, gold, tag, w, h, spp, depth):
    name, cq = tag.rsplit("_cq", 1)
    s = oracle.load_obj(os.path.join(SCENES, name + ".obj"), int(cq))
    lin, rgb, stats = oracle.render(s, w, h, spp, depth)
    key = f"{tag}_{w}x{h}_s{spp}_d{depth}"
    assert np.array_equal(bits(lin), bits(gold[key + "_linear"]))
    assert np.array_equal(rgb, gold[key + "_rgb8"])
    assert stats[1] == w * h * spp and stats[0] >= stats[1]
    oracle.free(s)


def test_rng_stream_persists_across_frames(oracle, gold):
    for tag in ("cbox_cq0", "cbox_quads_cq0"):
        name = tag[:-4]
        s = oracle.load_obj(os.path.join(SCENES, name + ".obj"))
        rng = oracle.render_init(32, 32)
        oracle.render(s, 32, 32, 4, 5, rng=rng)
        lin2, _, _ = oracle.render(s, 32, 32, 4, 5, rng=rng)
        assert np.array_equal(bits(lin2), bits(gold[f"{tag}_32x32_s4_d5_frame2_linear"]))
        oracle.free(s)


@pytest.mark.parametrize("name", ["cbox", "cbox_quads"])
def test_random_rays_bvh_and_linear(oracle, gold, name):
    s = oracle.load_obj(os.path.join(SCENES, name + ".obj"))
    rays = gold[name + "_rays"][:800]
    for mode, key in ((0, "bvh"), (1, "lin")):
        ids, ts = oracle.intersect(s, rays, mode=mode)
        assert np.array_equal(ids, gold[f"{name}_rays_{key}_ids"][:800].astype(np.int32))
        assert np.array_equal(bits(ts), bits(gold[f"{name}_rays_{key}_t"][:800]))
    oracle.free(s)


def test_tonemap_matches_reference_bytes(oracle, gold):
    lin = gold["cbox_cq0_48x48_s8_d8_linear"]
    assert np.array_equal(oracle.tonemap(lin), gold["cbox_cq0_48x48_s8_d8_rgb8"])


def test_segments_per_path_statistics(oracle):
    """2.730 segments/path at depth 5 on the Cornell box (128x128x8)."""
    s = oracle.load_obj(os.path.join(SCENES, "cbox.obj"))
    _, _, st = oracle.render(s, 128, 128, 8, 5)
    assert abs(st[0] / st[1] - 2.730) < 0.02
    assert abs(st[2] / st[0] - 14.0) < 0.3 and abs(st[3] / st[0] - 15.5) < 0.3    # reference BVH2 work per segment
    oracle.free(s)


@pytest.mark.parametrize("tag", ["cbox_cq0", "cbox_quads_cq0", "cbox_quads_cq1"])
def test_oracle_occluded_and_visibility_match_reference(oracle, tag):
    """Scene::occluded (scene.h:132-190) and render_radiosity (integrator.h:460-504) restated in
    oracle/pt_oracle.c against the reference's own code (tests/golden/f4_vectors.npz, make_golden_f4.py): the any-hit
    booleans, the visibility image and the RNG states it leaves behind must all be identical."""
    g = np.load(os.path.join(GOLDEN, "f4_vectors.npz"))
    name, cq = tag.rsplit("_cq", 1)
    s = oracle.load_obj(os.path.join(SCENES, name + ".obj"), convert_quads=bool(int(cq)))
    assert np.array_equal(oracle.occluded(s, g["segs"], mode=0), g[tag + "_occ"])
    assert np.array_equal(oracle.occluded(s, g["segs"], mode=1), g[tag + "_occ_lin"])
    for (w, h, spp) in ((48, 36, 3), (64, 40, 1)):
        rgb, rng = oracle.visibility(s, w, h, spp, radiosity=g[tag + "_rad"])
        assert np.array_equal(rgb, g[f"{tag}_vis_{w}x{h}_s{spp}"])
        st = np.frombuffer(rng, np.uint32).reshape(h * w, -1)[:, :6]
        assert np.array_equal(st, g[f"{tag}_visrng_{w}x{h}_s{spp}"])
    oracle.free(s)


# ---- grid-guided / MIS sampling: oracle pinned to the reference's integrator + Grid ----------
GUIDED = json.load(open(os.path.join(GOLDEN, "guided_vectors.json")))


def synthetic_grids(n_prims):
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_guided", os.path.join(GOLDEN, "make_golden_guided.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.synthetic_grids(n_prims)


@pytest.mark.parametrize("tag", sorted(GUIDED))
def test_guided_sampling_matches_reference_bit_for_bit(oracle, tag):
    """pto_render_guided (Grid::sample / computePDF / sampleMIS / the mode switch of integrator()) against the reference's
    own code run on the host (tests/golden/make_golden_guided.py): radiance AND final RNG states identical bit for bit,
    i.e. every draw, every branch and every double promotion of M_PI matches."""
    m = GUIDED[tag]
    gold = np.load(os.path.join(GOLDEN, "guided_vectors.npz"))
    s = oracle.load_obj(os.path.join(SCENES, m["scene"]), m["convert_quads"])
    oracle.set_grids(s, synthetic_grids(m["n_prims"]), m["bsdf_fraction"])
    lin, stats, rng = oracle.render_guided(s, m["w"], m["h"], m["spp"], m["depth"], m["mode"])
    st = np.frombuffer(bytes(rng), np.uint32).reshape(-1, 6)
    assert np.array_equal(st, gold[tag + "_rng"])
    assert np.array_equal(bits(lin), bits(gold[tag + "_linear"]))
    assert stats[1] == m["w"] * m["h"] * m["spp"]
    oracle.free(s)


def test_guided_modes_without_grids_equal_bsdf_sampling(oracle):
    """No radiosity data => initGridFromPrimitive finds all-zero grids => cosine fallback (integrator.h:258-261): every
    mode reproduces SAMPLING_BSDF exactly."""
    s = oracle.load_obj(os.path.join(SCENES, "cbox.obj"), 0)
    ref, _, _ = oracle.render(s, 24, 20, 4, 5)
    for mode in (1, 3):
        lin, _, _ = oracle.render_guided(s, 24, 20, 4, 5, mode)
        assert np.array_equal(bits(lin), bits(ref))
    oracle.set_grids(s, np.zeros((oracle.num_prims(s) if hasattr(oracle, "num_prims") else 32, 256, 3), np.float32))
    lin, _, _ = oracle.render_guided(s, 24, 20, 4, 5, 3)
    assert np.array_equal(bits(lin), bits(ref))
    oracle.free(s)


# ---- a scene with a real tree: 2 566 primitives, goldens from the reference's host build (tests/golden/make_golden_tree.py) ----
@pytest.fixture(scope="module")
def tree(pkg, tmp_path_factory):
    from cuda_pathtracer_b200 import scenegen
    g = np.load(os.path.join(GOLDEN, "tree_vectors.npz"))
    path = os.path.join(tmp_path_factory.mktemp("tree"), "tree.obj")
    n = scenegen.write_tree_obj(path, pkg.parse_scene_file(os.path.join(SCENES, "cbox_quads.obj")), level=int(g["level"]))
    assert n == len(g["prims"]) == 2566
    return path, g


def test_tree_scene_oracle_pinned_on_a_deep_bvh(oracle, tree):
    """The oracle's loadOBJ, BVHBuilder (median split, padded boxes), Scene::intersect through the BVH2 AND the linear scan, and the
    integrator on a 1 681-node tree: bit for bit what the reference's own code returns, including the 11 of 6 000 adversarial rays
    where the reference's BVH and its linear scan disagree with each other."""
    path, g = tree
    s = oracle.load_obj(path)
    prims, ref = oracle.prims(s), g["prims"]
    assert np.array_equal(prims["type"], ref[:, 0].view(np.int32))
    tri = prims["type"] == 0
    v = prims["v"].reshape(len(prims), 12)
    assert np.array_equal(bits(v[tri, :9]), bits(ref[tri, 1:10])) and np.array_equal(bits(v[~tri]), bits(ref[~tri, 1:13]))
    assert np.array_equal(bits(prims["n"]), bits(ref[:, 13:16])) and np.array_equal(bits(prims["bsdf"]), bits(ref[:, 16:19]))
    assert len(oracle.nodes(s)) == 1681
    ids, ts = oracle.primary_ids(s, 160, 160)
    assert np.array_equal(ids, g["ids"]) and np.array_equal(bits(ts), bits(g["ids_t"]))
    rays = g["rays"][:2500]
    for mode, key in ((0, "bvh"), (1, "lin")):
        i, t = oracle.intersect(s, rays, mode=mode)
        assert np.array_equal(i, g["rays_" + key][:2500])
        hit = i >= 0
        assert np.array_equal(bits(t[hit]), bits(g["rays_" + key + "_t"][:2500][hit]))
    assert (g["rays_bvh"] != g["rays_lin"]).sum() == 11
    lin, _, stats = oracle.render(s, 48, 48, 4, 8)
    assert np.array_equal(bits(lin), bits(g["render_48x48_4spp_d8"]))
    oracle.free(s)
