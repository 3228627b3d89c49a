"""ptx_trace_rays_noalloc (include/ptx.h): the allocation-free variant of the closest-hit query, whose device staging is a
file-scope device array of the library instead of per-call cudaMalloc.  It has no counterpart in the reference;
its contract is "the same answer as ptx_trace_rays", checked here bit for bit -- ids and
distances are integer / bit-exact work in every other test of this suite -- on two calls in a row (the second smaller than the
first, so stale contents of the shared array would show), on both scenes, and at the size limit.

Collected last on purpose (file name).
"""
import os

import numpy as np
import pytest

from conftest import SCENES

pytestmark = pytest.mark.gpu


def random_rays(rng, n):
    o = rng.uniform(-2.5, 2.5, (n, 3)) + np.array([0.0, 2.5, -3.0])
    d = rng.normal(size=(n, 3))
    return np.concatenate([o, d], 1).astype(np.float32)


@pytest.mark.parametrize("name", ["cbox.obj", "cbox_quads.obj"])
def test_noalloc_trace_matches_trace_rays_twice(pkg, ctx, name):
    scene = ctx.load_scene(os.path.join(SCENES, name))
    rng = np.random.default_rng(11)
    for n in (50_000, 1_237):                                   # two calls in a row through the C ABI, the second one smaller
        rays = random_rays(rng, n)
        ids_ref, t_ref = ctx.trace_rays(scene, rays)
        ids, t = ctx.trace_rays_noalloc(scene, rays)
        assert (ids_ref >= 0).mean() > 0.5
        assert np.array_equal(ids, ids_ref)
        assert np.array_equal(t.view(np.uint32), t_ref.view(np.uint32))
    scene.close()


def test_noalloc_limit_and_empty(pkg, ctx):
    scene = ctx.load_scene(os.path.join(SCENES, "cbox.obj"))
    rng = np.random.default_rng(12)
    rays = random_rays(rng, pkg.NOALLOC_MAX_RAYS)               # exactly the maximum
    ids_ref, t_ref = ctx.trace_rays(scene, rays)
    ids, t = ctx.trace_rays_noalloc(scene, rays)
    assert np.array_equal(ids, ids_ref) and np.array_equal(t.view(np.uint32), t_ref.view(np.uint32))
    with pytest.raises(pkg.PtxError) as e:                      # one more is rejected, with a clear message, before any device work
        ctx.trace_rays_noalloc(scene, random_rays(rng, pkg.NOALLOC_MAX_RAYS + 1))
    assert "PTX_NOALLOC_MAX_RAYS" in str(e.value)
    ids, t = ctx.trace_rays_noalloc(scene, np.zeros((0, 6), np.float32))
    assert len(ids) == 0 and len(t) == 0
    ids2, t2 = ctx.trace_rays_noalloc(scene, rays[:100])        # still fine after the rejected call
    assert np.array_equal(ids2, ids_ref[:100]) and np.array_equal(t2.view(np.uint32), t_ref[:100].view(np.uint32))
    scene.close()
