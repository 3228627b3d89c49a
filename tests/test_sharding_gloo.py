"""Multi-process (world_size 2 and 3, gloo, CPU) tests of the host side of multi-GPU rendering: the ranks' shares
are disjoint and complete, and the film combine (one all-reduce of the accumulation buffer) reassembles the frame.
Device rendering per shard is covered by the -m gpu tests (test_tile_sharding_..., test_philox_image_independent_...)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, w, h, spp, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    pkg = ge.load_package()
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        npix = w * h
        truth = (np.arange(npix, dtype=np.float64) * 0.001 + 1.0)          # "radiance" of pixel p for one sample
        # --- tiles: every rank owns whole pixels; film combine = sum of disjoint films
        own = pkg.shard_owned_pixels(w, h, pkg.SHARD_TILES, rank, world)
        film = np.zeros((npix, 4), np.float32)
        film[own, 0] = (truth[own] * spp).astype(np.float32)
        film[own, 3] = 1.0                                                    # ownership marker
        t = torch.from_numpy(film)
        dist.all_reduce(t)
        ok_tiles = bool(np.all(t.numpy()[:, 3] == 1.0)) and np.allclose(t.numpy()[:, 0], truth * spp, rtol=1e-6)
        # --- sample ranges: every rank renders all pixels for its samples
        first, count = pkg.shard_sample_range(spp, pkg.SHARD_SAMPLES, rank, world)
        film2 = np.zeros((npix, 4), np.float32)
        film2[:, 0] = (truth * count).astype(np.float32); film2[:, 3] = count
        t2 = torch.from_numpy(film2)
        dist.all_reduce(t2)
        ok_samples = bool(np.all(t2.numpy()[:, 3] == spp)) and np.allclose(t2.numpy()[:, 0], truth * spp, rtol=1e-5)
        ranges = [None] * world
        dist.all_gather_object(ranges, (first, count))
        contiguous = all(ranges[i][0] + ranges[i][1] == ranges[i + 1][0] for i in range(world - 1)) and ranges[0][0] == 0 and \
            ranges[-1][0] + ranges[-1][1] == spp
        counts = [None] * world
        dist.all_gather_object(counts, int(len(own)))
        q.put((rank, ok_tiles, ok_samples, contiguous, sum(counts) == npix))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,spp", [(2, 50, 23, 7), (3, 64, 36, 16), (2, 17, 3, 1)])
def test_sharding_film_combine_gloo(built, world, w, h, spp):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, w, h, spp, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_tiles, ok_samples, contiguous, complete in res:
        assert ok_tiles and ok_samples and contiguous and complete, (rank, ok_tiles, ok_samples, contiguous, complete)


def test_shard_arithmetic_single_process(pkg):
    for (w, h, world) in [(50, 23, 2), (64, 36, 3), (1920, 1080, 8), (7, 5, 4), (3, 2, 8)]:
        seen = np.zeros(w * h, np.int32)
        for r in range(world):
            own = pkg.shard_owned_pixels(w, h, pkg.SHARD_TILES, r, world)
            assert own.max(initial=0) < w * h
            seen[own] += 1
        assert (seen == 1).all()
        assert np.array_equal(pkg.shard_owned_pixels(w, h, pkg.SHARD_NONE, 0, 1), np.arange(w * h, dtype=np.uint32))
    for spp, world in [(1024, 8), (7, 3), (1, 4), (0, 2)]:
        tot = 0
        for r in range(world):
            f, c = pkg.shard_sample_range(spp, pkg.SHARD_SAMPLES, r, world)
            assert f == tot
            tot += c
        assert tot == spp


@pytest.mark.parametrize("w,h,world", [(50, 23, 2), (64, 36, 3), (333, 77, 8), (1920, 1080, 8), (17, 4, 5)])
def test_combine_slice_layout_is_the_inverse_of_the_owner_map(pkg, w, h, world):
    """ptx_film_combine, tiles: rank r packs its owned pixels in slot order into slice r of the gather buffer (per_rank = what rank 0
    owns, the largest share), and k_film_unpack_all reads pixel p back from (owner, local slot) = (band % world, (band / world) * band_px
    + p - band * band_px) with bands of 4 rows.  That formula must invert ptx_shard_owned_pixels for every rank -- checked here on
    the host, the kernel itself runs in tests/test_gpu_combine.py and in every N > 1 bench run."""
    import numpy as np
    band_px = 4 * w
    per_rank = len(pkg.shard_owned_pixels(w, h, pkg.SHARD_TILES, 0, world))
    gathered = np.full(per_rank * world, -1, np.int64)
    for r in range(world):
        owned = pkg.shard_owned_pixels(w, h, pkg.SHARD_TILES, r, world)
        assert len(owned) <= per_rank
        gathered[r * per_rank: r * per_rank + len(owned)] = owned          # what k_film_pack_owned writes: slot i <- pixel owned[i]
    p = np.arange(w * h)
    band = p // band_px
    src = (band % world) * per_rank + (band // world) * band_px + (p - band * band_px)
    assert np.array_equal(gathered[src], p)
