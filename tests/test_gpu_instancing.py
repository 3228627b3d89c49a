"""Two-level instancing (ptx_load_opts.keep_instances; csrc/traverse_inst.h) against the flattened load of the same .pbrt file -- which is
what the reference's loadPBRT produces (utils/pbrt_loader.h:197, 282-344) and what tests/golden/pbrt_vectors.npz pins bit for bit.
Numbering: identical.  Hit distances: the instanced scene intersects in object space, so they agree to rounding, not bit for bit."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def scenes(pkg, ctx, tmp_path_factory):
    from cuda_pathtracer_b200 import scenegen
    path = str(tmp_path_factory.mktemp("inst") / "c5_small.pbrt")
    n = scenegen.write_c5_pbrt(path, instances=40, level=4, n_lights=9)
    flat = ctx.load_scene(path, mirror_from_mtl=True)
    inst = ctx.load_scene(path, mirror_from_mtl=True, keep_instances=True)
    side = 4; room = 4.0 * side
    cam_args = ((room / 2, room / 2, room * 1.9), (room / 2, room / 2, 0.0), (0.0, 1.0, 0.0), 40.0)
    yield flat, inst, n, cam_args
    flat.close(); inst.close()


def test_instanced_scene_layout(pkg, ctx, scenes):
    flat, inst, n, _ = scenes
    sf, si = flat.stats(), inst.stats()
    assert sf["n_instances"] == 0 and si["n_instances"] == 41                  # 40 instances + the world's own shapes under the identity
    assert sf["n_triangles"] == si["n_triangles"] == n                         # the same flattened primitive count
    assert si["stored_triangles"] == 2 * 20 * 4 ** 4 + 10 + 2 * 9              # two object definitions + room + lights, each ONCE
    assert si["triangle_bytes"] < sf["triangle_bytes"] / 10


def test_instanced_hits_equal_flattened(pkg, ctx, scenes):
    flat, inst, n, cam_args = scenes
    w = h = 384
    cam = pkg.look_at_camera(*cam_args, w, h)
    idf, tf = ctx.trace_primary_ids(flat, cam, w, h)
    idi, ti = ctx.trace_primary_ids(inst, cam, w, h)
    assert (idf >= 0).mean() > 0.9
    same = idf == idi
    assert same.mean() >= 0.9995, float(same.mean())                           # differences only where two hits tie to rounding
    assert np.allclose(ti[same], tf[same], rtol=2e-5, atol=1e-5)
    rng = np.random.default_rng(4)
    o = rng.uniform(1.0, 15.0, (20000, 3)); d = rng.normal(size=(20000, 3))
    rays = np.concatenate([o, d], 1).astype(np.float32)
    idf, tf = ctx.trace_rays(flat, rays)
    idi, ti = ctx.trace_rays(inst, rays)
    same = idf == idi
    assert same.mean() >= 0.9995, float(same.mean())
    close = np.isclose(ti[same], tf[same], rtol=1e-4, atol=1e-4)               # grazing hits on a scaled / rotated instance: object-space
    assert close.mean() >= 0.999, float(close.mean())                          # arithmetic moves t by more than a few ulps there
    assert idi.max() < n


@pytest.mark.parametrize("rng_mode", [0, 1])
def test_instanced_render_close_to_flattened(pkg, ctx, scenes, rng_mode):
    flat, inst, n, cam_args = scenes
    w, h, spp, depth = 160, 120, 16, 8
    cam = pkg.look_at_camera(*cam_args, w, h)
    film = ctx.create_film(w, h)
    stf = ctx.render(flat, cam, film, spp=spp, max_depth=depth, rng_mode=rng_mode)
    a = film.linear()
    film.reset()
    sti = ctx.render(inst, cam, film, spp=spp, max_depth=depth, rng_mode=rng_mode)
    b = film.linear()
    film.close()
    assert sti["paths"] == stf["paths"] == w * h * spp
    assert abs(int(sti["segments"]) - int(stf["segments"])) <= 2e-3 * stf["segments"]
    # same streams, same decisions except where a last-bit difference of a hit distance flips one: most pixels agree closely
    # (a pixel where one path diverged moves by a whole sample at 16 spp -- up to the emitters' radiance 12 -- so a mean relative error
    #  is meaningless here: count the pixels that agree instead, and compare the image means)
    rel = np.abs(a.astype(np.float64) - b) / np.maximum(a, 1e-3)
    same = (rel <= 1e-4).all(-1)
    print(f"\n[instancing] rng {rng_mode}: pixels equal to 1e-4: {same.mean():.4f}, means {a.mean():.6f} / {b.mean():.6f}, segments {stf['segments']} / {sti['segments']}")
    assert same.mean() >= 0.9, float(same.mean())
    assert abs(float(a.mean()) - float(b.mean())) <= 2e-2 * float(a.mean())


def test_instanced_shadow_queries_and_visibility_view(pkg, ctx, scenes):
    """Scene::occluded and the render_radiosity view through the two-level structure (trace_any_instanced, k_visibility<INST>) and the
    allocation-free ray query: the same answers as the flattened scene, but for rays that graze within rounding."""
    flat, inst, n, cam_args = scenes
    rng = np.random.default_rng(12)
    segs = np.concatenate([rng.uniform(1.0, 15.0, (30000, 3)), rng.uniform(1.0, 15.0, (30000, 3))], 1).astype(np.float32)
    of, oi = ctx.trace_occluded(flat, segs), ctx.trace_occluded(inst, segs)
    assert 0.2 < of.mean() < 0.98
    assert (of != oi).mean() < 5e-4, float((of != oi).mean())
    w, h, spp = 160, 120, 4
    cam = pkg.look_at_camera(*cam_args, w, h)
    film = ctx.create_film(w, h)
    rad = rng.uniform(0.0, 0.6, (n, 3)).astype(np.float32)
    out = []
    for sc in (flat, inst):
        film.reset(2023)
        img = ctx.render_visibility(sc, cam, film, spp, radiosity=rad)
        out.append((img, film.rng_state()))
    assert np.array_equal(out[0][1], out[1][1])                                 # the same streams, the same number of draws
    differ = (out[0][0] != out[1][0]).any(-1).mean()
    assert differ < 5e-3, float(differ)                                          # a pixel differs where one of its 4 rays ties to rounding
    assert out[0][0].mean() > 20
    film.close()
    rays = np.concatenate([rng.uniform(1.0, 15.0, (5000, 3)), rng.normal(size=(5000, 3))], 1).astype(np.float32)
    ia, ta = ctx.trace_rays(inst, rays)
    ib, tb = ctx.trace_rays_noalloc(inst, rays)
    assert np.array_equal(ia, ib) and np.array_equal(ta.view(np.uint32), tb.view(np.uint32))


def test_instanced_scene_unsupported_calls(pkg, ctx, scenes):
    flat, inst, n, cam_args = scenes
    with pytest.raises(pkg.PtxError):
        inst.rebuild(builder=0)
    cam = pkg.look_at_camera(*cam_args, 32, 32)
    film = ctx.create_film(32, 32)
    with pytest.raises(pkg.PtxError):
        ctx.render(inst, cam, film, spp=1, pipeline=pkg.PIPELINE_WAVEFRONT)
    film.close()
