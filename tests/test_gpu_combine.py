"""ptx_comm_* / ptx_film_combine / PTX_RENDER_ASYNC through the C ABI on ONE GPU (a communicator of one rank): the pack ->
ncclAllGather -> unpack path and the in-place ncclAllReduce path must leave the film exactly as a plain render does, and an
asynchronous render must report the same statistics once they are asked for.  The multi-rank behaviour is checked by
bench.py itself on every N > 1 run (`multi_rank_check`: combined film == unsharded film, asserted) and by tools/combine_check.py."""
import os

import numpy as np
import pytest

from conftest import SCENES

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def comm(pkg, ctx):
    try:
        c = ctx.create_comm(pkg.comm_unique_id(), 0, 1)
    except pkg.PtxError as e:
        if e.code == -5:
            pytest.skip(f"NCCL not loadable here: {e}")
        raise
    yield c
    c.close()


@pytest.mark.parametrize("w,h", [(96, 54), (50, 23)])          # 23 rows: the last band of 4 rows is partial
def test_combine_single_rank_is_identity(pkg, ctx, comm, w, h):
    scene = ctx.load_scene(os.path.join(SCENES, "cbox_quads.obj"))
    cam = pkg.default_camera(w, h)
    film = ctx.create_film(w, h)
    st0 = ctx.render(scene, cam, film, spp=8, max_depth=8)
    plain = film.linear()
    film.reset()
    assert ctx.render(scene, cam, film, spp=8, max_depth=8, shard_mode=pkg.SHARD_TILES, shard_rank=0, shard_world=1, async_=True) is None
    film.combine(comm, pkg.SHARD_TILES)                         # no host synchronisation up to here
    assert np.array_equal(film.linear(), plain)
    st = ctx.render_stats()                                     # completes the asynchronous render's statistics
    assert st["segments"] == st0["segments"] and st["paths"] == st0["paths"] and st["render_ms"] > 0
    # counter RNG, sample ranges
    ctx.render(scene, cam, film, spp=8, max_depth=8, rng_mode=pkg.RNG_PHILOX)
    plain = film.linear()
    ctx.render(scene, cam, film, spp=8, max_depth=8, rng_mode=pkg.RNG_PHILOX, shard_mode=pkg.SHARD_SAMPLES, shard_rank=0, shard_world=1, async_=True)
    film.combine(comm, pkg.SHARD_SAMPLES)
    assert np.allclose(film.linear(), plain, rtol=1e-6, atol=1e-7)
    with pytest.raises(pkg.PtxError):
        film.combine(comm, pkg.SHARD_NONE)
    film.close(); scene.close()


def test_async_render_then_sync_render(pkg, ctx):
    """An asynchronous render followed at once by another render: the second call first completes the first one."""
    scene = ctx.load_scene(os.path.join(SCENES, "cbox.obj"))
    cam = pkg.default_camera(64, 64)
    film = ctx.create_film(64, 64)
    ref = ctx.render(scene, cam, film, spp=4, max_depth=5)
    ref2 = ctx.render(scene, cam, film, spp=4, max_depth=5, accumulate=True)
    want = film.linear()
    film.reset()
    ctx.render(scene, cam, film, spp=4, max_depth=5, async_=True)
    st2 = ctx.render(scene, cam, film, spp=4, max_depth=5, accumulate=True)
    assert st2["segments"] == ref2["segments"]
    assert np.array_equal(film.linear(), want)
    film.close(); scene.close()


def test_cpp_multi_gpu_example_runs(built, tmp_path):
    """examples/multi_gpu.cpp (C++ threads + the C ABI only) on as many GPUs as this box has (1 or 2 here): the combined film equals the
    unsharded one bit for bit, every rank holds the same film, and rank 0 writes the PNG."""
    import subprocess
    pkg = built.load_package()
    exe = built.build_multi_gpu_example()
    n = 1
    try:                                  # a second GPU?  (not via torch: importing it AFTER this process loaded the system NCCL fails,
        pkg.Context(1).close(); n = 2     #  PyTorch bundles a newer NCCL under the same soname -- see the note at ptx_comm_create)
    except pkg.PtxError:
        pass
    png = tmp_path / "out.png"
    r = subprocess.run([exe, os.path.join(SCENES, "cbox_quads.obj"), "160", "90", "16", "8", str(n), str(png)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "IDENTICAL" in r.stdout
    assert png.exists() and png.stat().st_size > 1000
