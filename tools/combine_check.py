"""Multi-GPU check of the library's film combine (run under torchrun on N GPUs):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/combine_check.py
Every rank renders its share of C2 (64 spp) and C4-like frames, ptx_film_combine exchanges them, and every rank compares the film
it ends up with against the same frame rendered unsharded on its own GPU: bit for bit (tiles), 1e-5 (sample ranges)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import __graft_entry__ as ge
pkg = ge.load_package()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
if rank == 0:
    uid.copy_(torch.tensor(list(pkg.comm_unique_id()), dtype=torch.uint8))
dist.broadcast(uid, 0)
ctx = pkg.Context(local)
comm = ctx.create_comm(bytes(uid.cpu().tolist()), rank, world)
ok = True
for (w, h, spp) in ((1920, 1080, 64), (50, 23, 16), (333, 77, 8)):
    scene = ctx.load_scene(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
    cam = pkg.default_camera(w, h)
    film = ctx.create_film(w, h)
    for rng, mode in ((pkg.RNG_XORWOW_COMPAT, pkg.SHARD_TILES), (pkg.RNG_PHILOX, pkg.SHARD_SAMPLES)):
        film.reset()
        ctx.render(scene, cam, film, spp=spp, max_depth=16, rng_mode=rng)
        whole = film.linear()
        film.reset()
        ctx.render(scene, cam, film, spp=spp, max_depth=16, rng_mode=rng, shard_mode=mode, shard_rank=rank, shard_world=world, async_=True)
        film.combine(comm, mode)
        got = film.linear()
        nd = int((got != whole).any(-1).sum())
        rel = float(np.max(np.abs(got.astype(np.float64) - whole) / np.maximum(np.abs(whole), 1e-3)))
        good = nd == 0 if mode == pkg.SHARD_TILES else rel <= 1e-5
        ok &= good
        print(json.dumps(dict(rank=rank, world=world, w=w, h=h, spp=spp, mode=mode, pixels_differing=nd, max_rel=rel, ok=bool(good))), flush=True)
    film.close(); scene.close()
comm.close(); ctx.close()
flag = torch.tensor([0.0 if ok else 1.0], device="cuda"); dist.all_reduce(flag)
dist.destroy_process_group()
sys.exit(0 if flag.item() == 0 else 1)
