#!/usr/bin/env python
"""bench.py -- headline benchmark of the path-tracing hot path (contract: see the task statement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--rng compat|philox]
                    [--pipeline persist|loop|fused|wavefront] [--workload c1|c2|c3|c4]

One "step" = one frame of the hot path: ptx_render (raygen/extend/shade/compact) + film resolve.
N = 1 workload = BASELINE.json configs[1]: scenes/cbox_quads.obj, 1920x1080, 1024 spp, depth 16, reference
Russian roulette.  N > 1: the same frame, sharded over the ranks (tiles with the compat RNG, sample ranges
with the counter RNG) with the film combined by one NCCL all-reduce ("scaling": "strong").
Metric = Mpaths*bounces/s = 1e-6 * path segments (exact device counter) / seconds, whole job.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene, width, height, spp, depth)
    "c1": ("cbox.obj", 512, 512, 64, 8),
    "c2": ("cbox_quads.obj", 1920, 1080, 1024, 16),
    "c4": ("generated:c4", 3840, 2160, 4096, 8),        # mirror tall box + emissive short box (scenegen.c4_prims)
    # generated on the box (cuda-pathtracer_b200/scenegen.py): two displaced icospheres, 10 485 772 triangles, Cornell enclosure
    "c3": ("generated:c3_level9", 1920, 1080, 256, 8),
    # generated on the box as a .pbrt file (scenegen.write_c5_pbrt) and loaded through ptx_scene_load with keep_instances:
    # 305 instances of a 327 680-triangle mesh = 99 942 538 triangles, 64 emissive quads.  Not a default.
    "c5": ("generated:c5", 3840, 2160, 1024, 8),        # 305 instances x 327 680 triangles = 99.9 M, 64 emissive quads; kept TWO-LEVEL
}
METRIC = "Mpaths*bounces/s"


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.reasons, self.max_mhz, self.stop_flag = gpu_index, [], set(), None, False

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                r = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5)
                f = [x.strip() for x in r.stdout.strip().split(",")]
                self.samples.append(float(f[0])); self.max_mhz = float(f[1])
                for n, v in zip(names, f[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


def ref_host_run(scene, w, h, spp, depth, frames, threads=0):
    """threads = 0: every host core this process may use -- stated explicitly, because torchrun exports OMP_NUM_THREADS=1
    to its workers and the reference arm must not be slowed down by the launcher."""
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_host")
    if threads <= 0:
        threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = str(threads)
    r = subprocess.run([exe, "render", scene, str(w), str(h), str(spp), str(depth), str(frames), "-", "0", str(threads)],
                       capture_output=True, text=True, check=True, env=env)
    return json.loads(r.stdout.strip().splitlines()[-1])


def ref_cuda_run(scene, w, h, spp, depth, frames=1, orig=False, gpu=0):
    """The reference's own CUDA kernels on this box's GPU (oracle/_ref/ref_cuda: its unmodified headers compiled for sm_100a;
    `render_twin` = `render` with max_depth as a parameter, orig=True = `render` exactly as it ships, depth 5).  CUDA-event ms."""
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_cuda")
    env = dict(os.environ)
    env["CUDA_VISIBLE_DEVICES"] = str(gpu)
    r = subprocess.run([exe, "render", scene, str(w), str(h), str(spp), str(depth), str(frames), "-", "0", "0", "1" if orig else "0"],
                       capture_output=True, text=True, check=True, env=env, timeout=900)
    return json.loads(r.stdout.strip().splitlines()[-1])


def gpu_baseline(scene_path, w, h, spp, depth, segments_per_frame, segments_depth5=None):
    """R-GPU baseline of BASELINE.json's metric: the reference's CUDA build on the same B200, same frame, right before the
    timed region.  The reference has no segment counter; its streams are the product's (compat RNG), so the product's exact
    count for the same frame converts its ms/frame into the metric."""
    out = {"unit": METRIC, "impl": "oracle/_ref/ref_cuda (reference headers, nvcc -O3 sm_100a, block 16x16)"}
    try:
        ref_cuda_run(scene_path, 256, 256, 4, depth)            # warm-up: context creation, module load
        j = ref_cuda_run(scene_path, w, h, spp, depth, frames=2)
        out.update(kernel="render_twin (render with max_depth as a parameter)", ms_per_frame=j["ms_per_frame"], ms_init=j["ms_init"],
                   value=segments_per_frame / (j["ms_per_frame"] * 1e-3) * 1e-6, segments_per_frame=segments_per_frame,
                   segments_source="device counter of this build on the same frame (same RNG streams)")
        o = ref_cuda_run(scene_path, w, h, spp, depth, frames=2, orig=True)
        out["unmodified_render_depth5"] = {"kernel": "render", "ms_per_frame": o["ms_per_frame"],
                                           "value": (segments_depth5 / (o["ms_per_frame"] * 1e-3) * 1e-6) if segments_depth5 else None}
    except Exception as e:       # noqa: BLE001
        out["error"] = str(e)[:300]
    return out


def segments_per_path_estimate(scene, w, h, depth):
    """Exact segment counts come from the oracle (same streams as the reference): used to convert the
    reference's ms/frame into the metric, since the reference has no segment counter."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding
    o = oracle_binding.Oracle()
    s = o.load_obj(scene)
    sw, sh = max(16, w // 8), max(9, h // 8)
    _, _, st = o.render(s, sw, sh, 4, depth)
    o.free(s)
    return st[0] / st[1]


def run_reference(args, scene_path, w, h, spp, depth):
    """--impl reference: the reference's own CPU build (oracle/_ref/ref_host: its unmodified headers compiled
    for the host, OpenMP over rows) on a bounded sample of the workload, all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sample_spp = 2
    seg_per_path = segments_per_path_estimate(scene_path, w, h, depth)
    for _ in range(args.warmup):
        ref_host_run(scene_path, w, h, 1, depth, 1)
    t_ms = []
    for _ in range(args.steps):
        j = ref_host_run(scene_path, w, h, sample_spp, depth, 1)
        t_ms.append(j["ms_per_frame"])
        threads = j["threads"]
    ms = sum(t_ms) / len(t_ms)
    value = w * h * sample_spp * seg_per_path / (ms * 1e-3) * 1e-6
    sample = f"{w}x{h}, {sample_spp} of {spp} spp per step, depth {depth}"
    emit(({
        "impl": "reference", "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload_desc, "sample": sample, "segments_per_path": seg_per_path,
                   "segments_per_path_source": "C oracle on a 1/8-scale frame at 4 spp (same streams as the reference), not counted on the timed frame"},
        "cpu_baseline": {"value": value, "unit": METRIC, "cores": threads, "kind": "reference", "sample": sample, "host_cores": cores},
        "e2e": {"value": value, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def build_scene(pkg, ctx, workload, rank, max_leaf=0):
    """Scene + camera of a workload, replicated on this rank's GPU (one-time, outside every timed region)."""
    scene_name, w, h, _, _ = WORKLOADS[workload]
    cam = None
    if scene_name.startswith("generated:"):
        from cuda_pathtracer_b200 import scenegen
        base = pkg.parse_scene_file(os.path.join(ROOT, "scenes", "cbox_quads.obj"))
        if scene_name.endswith("c5"):
            pbrt = os.path.join("/tmp", f"c5_rank{rank}.pbrt")
            scenegen.write_c5_pbrt(pbrt, instances=305, level=7)
            scene = ctx.load_scene(pbrt, mirror_from_mtl=True, keep_instances=True)
            side = 7; room = 4.0 * side                       # camera of the generated file (scenegen.write_c5_pbrt)
            cam = pkg.look_at_camera((room / 2, room / 2, room * 1.9), (room / 2, room / 2, 0.0), (0.0, 1.0, 0.0), 40.0, w, h)
        elif scene_name.endswith("c4"):
            scene = ctx.scene_from_primitives(scenegen.c4_prims(base), builder=0, max_leaf=max_leaf)
        else:
            level = int(scene_name.rsplit("level", 1)[1])
            tris, mat, mats = scenegen.c3_scene(base, level)
            scene = ctx.scene_from_triangles(tris, mat, mats, builder=0, max_leaf=max_leaf)
            del tris, mat
    else:
        scene = ctx.load_scene(os.path.join(ROOT, "scenes", scene_name))
        if max_leaf:
            scene.rebuild(builder=0, max_leaf=max_leaf)
    return scene, (cam or pkg.default_camera(w, h))


class Run:
    """One workload on this rank: frames through the C ABI, sharded over `world` ranks, film combined by the LIBRARY
    (ptx_film_combine: NCCL on the context's stream); torch.distributed only carries barriers and the metric reductions."""

    def __init__(self, pkg, ctx, comm, dist, torch, rank, world, workload, spp, depth, rng, pipeline, args):
        import numpy as np
        self.np, self.pkg, self.ctx, self.comm, self.dist, self.torch, self.rank, self.world = np, pkg, ctx, comm, dist, torch, rank, world
        self.workload, self.spp, self.depth = workload, spp, depth
        _, self.w, self.h, _, _ = WORKLOADS[workload]
        self.rng_mode = pkg.RNG_PHILOX if rng == "philox" else pkg.RNG_XORWOW_COMPAT
        self.shard_mode = pkg.SHARD_NONE if world == 1 else (pkg.SHARD_SAMPLES if self.rng_mode == pkg.RNG_PHILOX else pkg.SHARD_TILES)
        self.scene, self.cam = build_scene(pkg, ctx, workload, rank, args.max_leaf)
        self.default_cam = not WORKLOADS[workload][0].endswith("c5")
        self.film = ctx.create_film(self.w, self.h)
        self.rgb_host = np.empty((self.h, self.w, 3), np.uint8)
        self.kw = dict(spp=spp, max_depth=depth, rng_mode=self.rng_mode, pipeline=pipeline, shard_mode=self.shard_mode, shard_rank=rank, shard_world=world,
                       loop_min_blocks=args.min_blocks, chunks=args.chunks, shade_batch=args.shade_batch,
                       cost_order=not getattr(args, "no_cost_order", False), bounds_cull=not getattr(args, "no_cull", False))

    def barrier(self):
        self.ctx.sync()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def step(self, host_output, stage_timing=False):
        """One frame.  host_output=False: the RGB8 result stays in HBM. True: it is copied to the host (renderFrame, application.h:211-212).
        Nothing synchronises the host between render and combine (PTX_RENDER_ASYNC); the resolve at the end of the frame does."""
        self.ctx.flush_l2()                           # 256 MB memset on the render stream: nothing of the last frame stays in L2
        multi = self.world > 1
        st = self.ctx.render(self.scene, self.cam, self.film, stage_timing=stage_timing, async_=multi and not stage_timing, **self.kw)
        if multi:
            self.film.combine(self.comm, self.shard_mode)
        if host_output:
            self.film.rgb8(out=self.rgb_host)
        else:
            self.film.resolve_device()
        return st if st is not None else self.ctx.render_stats()

    def measure(self, steps, warmup, e2e=True):
        pkg = self.pkg
        for _ in range(warmup):
            self.step(False)
        # ---- timed region A: inputs resident in HBM, device events on the launching stream
        self.barrier()
        self.ctx.timer_start()
        seg = launches = culled = 0
        for _ in range(steps):
            st = self.step(False)
            culled += st.get("culled_paths", 0)
            seg += st["segments"]; launches += st["kernel_launches"] + 1 + (3 if self.shard_mode == pkg.SHARD_TILES else (1 if self.world > 1 else 0))
        ms_dev = self.ctx.timer_stop()
        self.barrier()
        # ---- timed region B: end to end through the C ABI with host buffers
        seg_e2e, ms_e2e, culled_e2e = 0, 0.0, 0
        if e2e:
            self.barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                if self.default_cam:
                    self.cam = pkg.default_camera(self.w, self.h)    # host camera update + by-value upload, like renderFrame (application.h:161-164)
                st = self.step(True)
                seg_e2e += st["segments"]; culled_e2e += st.get("culled_paths", 0)
            self.ctx.sync()
            ms_e2e = (time.perf_counter() - t0) * 1e3
            self.barrier()
        out = dict(seg=seg, seg_e2e=seg_e2e, ms_dev=ms_dev, ms_e2e=ms_e2e, launches=launches, culled=culled)
        if self.world > 1:
            t = self.torch.tensor([float(seg), float(seg_e2e), float(culled), float(culled_e2e)], dtype=self.torch.float64, device="cuda")
            self.dist.all_reduce(t)
            out.update(culled=t[2].item())
            culled_e2e = t[3].item()
            m = self.torch.tensor([ms_dev, ms_e2e], dtype=self.torch.float64, device="cuda")
            self.dist.all_reduce(m, op=self.dist.ReduceOp.MAX)
            out.update(seg=t[0].item(), seg_e2e=t[1].item(), ms_dev=m[0].item(), ms_e2e=m[1].item())
        out["value"] = out["seg"] / (out["ms_dev"] * 1e-3) * 1e-6
        out["e2e_value"] = out["seg_e2e"] / (out["ms_e2e"] * 1e-3) * 1e-6 if e2e else None
        out["ms_per_step"] = out["ms_dev"] / steps
        # whole-pixel culling (ptx_render): the frame's paths whose every ray misses the scene's bounds are resolved without tracing;
        # `value` counts the WORKLOAD's segments (what the reference traces for this frame), value_traced only the traced ones
        out["value_traced"] = (out["seg"] - out["culled"]) / (out["ms_dev"] * 1e-3) * 1e-6
        out["e2e_value_traced"] = (out["seg_e2e"] - culled_e2e) / (out["ms_e2e"] * 1e-3) * 1e-6 if e2e else None
        out["culled_per_step"] = out["culled"] / steps
        out["seg_per_step"] = out["seg"] / steps
        return out

    def check_against_unsharded(self):
        """Multi-rank check, outside every timed region: every rank compares the combined film it holds with the same frame
        rendered unsharded on its own GPU -- bit for bit with tiles + compat RNG, to 1e-5 with sample ranges + counter RNG."""
        np, pkg = self.np, self.pkg
        self.film.reset()
        self.step(False)
        combined = self.film.linear()
        kw = dict(self.kw); kw.update(shard_mode=pkg.SHARD_NONE, shard_rank=0, shard_world=1)
        self.film.reset()
        self.ctx.render(self.scene, self.cam, self.film, **kw)
        whole = self.film.linear()
        self.film.reset()
        differing = int((combined != whole).any(-1).sum())
        relerr = np.abs(combined.astype(np.float64) - whole) / np.maximum(np.abs(whole), 1e-3)
        rel, mean_rel = float(relerr.max()), float(relerr.mean())
        # sample ranges: the same samples, added with float atomics in another order (and in two partial sums): mean <= 1e-6, worst pixel <= 1e-4
        ok = differing == 0 if self.shard_mode == pkg.SHARD_TILES else (mean_rel <= 1e-6 and rel <= 1e-4)
        res = {"mode": {1: "tiles: bit-exact", 2: "sample ranges: mean <= 1e-6, max <= 1e-4 relative (float atomics in another order)"}[self.shard_mode],
               "pixels_differing": differing, "max_rel_diff": rel, "mean_rel_diff": mean_rel, "ok": bool(ok), "rank": self.rank}
        if self.world > 1:
            flags = self.torch.tensor([0.0 if ok else 1.0], dtype=self.torch.float64, device="cuda")
            self.dist.all_reduce(flags)
            res["ranks_failing"] = int(flags[0].item())
            res["ok"] = res["ranks_failing"] == 0
        return res

    def close(self):
        self.film.close(); self.scene.close()


def roofline_of(run, m, st_t, args, clocks, sm_count):
    """roofline object of the dominant kernel of `run` (k_path_flat / k_path_persist: one launch per frame = the whole render).
    Large scenes (BVH8 + records > L2): HBM bound on node + triangle fetches, algorithmic bytes from the instrumented extend kernel.
    Cache-resident scenes: HBM is not the bound (the kernel moves < 1 byte per segment); the binding resource is instruction
    ISSUE: achieved = segments/s x warp-instructions per segment (from the committed ncu capture of this kernel, profiles/traffic.json),
    peak = SMs x 4 schedulers x the SM clock sampled during this run.  The HBM figures stay as a side note."""
    pkg, ctx = run.pkg, run.ctx
    peak, peak_src = read_peaks()
    scene_stats = run.scene.stats()
    k_ms = st_t["shade_ms"]
    is_flat = scene_stats.get("flat_units", 0) > 0 and args.pipeline in ("auto", "flat")
    kname = ("k_path_flat<rng> (one launch per frame: flat unit-list extend + shade + regenerate, all state in registers)" if is_flat else
             "k_path_persist<rng> (one launch per frame: BVH8 extend + shade + regenerate, all state in registers)")
    n_px = run.w * run.h / run.world
    n_chunks = compat_chunk_count(run.spp, args.chunks, n_px / (sm_count * 768.0))
    # per owned pixel and sample chunk: pixel id 4 + RNG state 24 + 24 + pixel sum 16 + 16 + progress flag 4 + 4; counter RNG: pixel id 4 + three float atomics 12 per path
    per_item = 92.0 * n_chunks * n_px if run.rng_mode == pkg.RNG_XORWOW_COMPAT else 16.0 * st_t["paths"]
    bytes_per_seg = per_item / max(st_t["segments"] - st_t.get("culled_paths", 0), 1)
    traced = st_t["segments"] - st_t.get("culled_paths", 0)          # the roofline is about what the kernel executes: traced segments only
    seg_per_s = traced / (k_ms * 1e-3) if k_ms > 0 else 0.0
    tkey = f"{run.workload}/{'philox' if run.rng_mode == pkg.RNG_PHILOX else 'compat'}/{'flat' if is_flat else 'persist'}"
    ent = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        ent = json.load(open(tpath)).get(tkey)
    common = {"kernel": kname, "kernel_ms_per_step": k_ms, "kernel_share_of_step": k_ms / m["ms_per_step"] if m["ms_per_step"] else None,
              "traffic": ent["bytes"] if ent and run.world == 1 and args.spp == 0 else None, "traffic_source": ent["source"] if ent else None}
    if scene_stats["node_bytes"] + scene_stats["triangle_bytes"] > 126e6:
        kwc = dict(run.kw); kwc.update(spp=2, pipeline=pkg.PIPELINE_WAVEFRONT, shard_mode=pkg.SHARD_NONE, shard_rank=0, shard_world=1)
        st_c = ctx.render(run.scene, run.cam, run.film, count_work=True, **kwc)
        n_node, n_tri = st_c["node_visits"] / max(st_c["segments"], 1), st_c["tri_tests"] / max(st_c["segments"], 1)
        bps = n_node * 80 + n_tri * 48
        # the instrumented run traces EVERY segment of the frame; the timed kernel does not trace the culled ones (whole-pixel culling),
        # each of which would have cost exactly one node visit (the root, all children missed): take those 80 bytes out
        f_culled = st_t.get("culled_paths", 0) / max(st_t["segments"], 1)
        all_per_s = st_t["segments"] / (k_ms * 1e-3) if k_ms > 0 else 0.0
        bps_exec = bps - f_culled * 80.0
        achieved = all_per_s * bps_exec * 1e-9
        return dict(common, bound="hbm", achieved=achieved, peak=peak, unit="GB/s", frac=achieved / peak, peak_source=peak_src, bytes_per_segment=bps,
                    bytes_per_segment_executed=bps_exec, culled_fraction_of_segments=f_culled,
                    traversal={"nodes_per_segment": n_node, "tris_per_segment": n_tri, "node_bytes": 80, "tri_bytes": 48})
    hbm = {"achieved": seg_per_s * bytes_per_seg * 1e-9, "peak": peak, "unit": "GB/s", "frac": seg_per_s * bytes_per_seg * 1e-9 / peak, "peak_source": peak_src,
           "bytes_per_segment": bytes_per_seg, "note": "algorithmic HBM bytes of a kernel that keeps all path state in registers: not the bound"}
    issue = (ent or {}).get("issue")
    if issue and clocks.get("sm_mhz"):
        wips = issue["warp_instructions_per_segment"]
        peak_issue = sm_count * 4 * clocks["sm_mhz"] * 1e6 * 1e-9            # G warp-instructions / s
        ach = seg_per_s * wips * 1e-9
        return dict(common, bound="issue", achieved=ach, peak=peak_issue, unit="G warp-instr/s", frac=ach / peak_issue,
                    peak_source=f"{sm_count} SMs x 4 schedulers x {clocks['sm_mhz']:.0f} MHz (nvidia-smi, median during the timed region)",
                    warp_instructions_per_segment=wips, lane_efficiency=issue.get("lane_efficiency"), issue_source=issue.get("source"), hbm=hbm)
    return dict(common, bound="hbm", **hbm)


def run_ours(args, w, h, spp, depth):
    import __graft_entry__ as ge
    pkg = ge.load_package()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    pipeline = {"fused": pkg.PIPELINE_FUSED, "wavefront": pkg.PIPELINE_WAVEFRONT, "loop": pkg.PIPELINE_LOOP, "persist": pkg.PIPELINE_PERSIST, "cta": pkg.PIPELINE_CTA,
                "flat": pkg.PIPELINE_FLAT, "auto": pkg.PIPELINE_DEFAULT}[args.pipeline]
    dist = torch = comm = None
    ctx = pkg.Context(local_rank)
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        # the library's own communicator: rank 0 makes the id, torch.distributed only ships its 128 bytes
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.tensor(list(pkg.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        comm = ctx.create_comm(bytes(uid.cpu().tolist()), rank, world)
    sm_count = ctx.device_info()["sm_count"]
    generated = WORKLOADS[args.workload][0].startswith("generated:")
    scene_path = os.path.join(ROOT, "scenes", WORKLOADS[args.workload][0])

    run = Run(pkg, ctx, comm, dist, torch, rank, world, args.workload, spp, depth, args.rng, pipeline, args)
    scene_stats = run.scene.stats()
    gpu_base = None
    if rank == 0 and not generated and not args.no_gpu_baseline:
        # one frame of ours for the exact segment counts (depth of the config, and depth 5 for the unmodified kernel), then the
        # reference's CUDA kernels on this very GPU; the other ranks wait at the first barrier of measure()
        kw1 = dict(run.kw); kw1.update(shard_mode=pkg.SHARD_NONE, shard_rank=0, shard_world=1)
        seg_full = ctx.render(run.scene, run.cam, run.film, **kw1)["segments"]
        kw5 = dict(kw1); kw5.update(max_depth=5)
        seg_d5 = ctx.render(run.scene, run.cam, run.film, **kw5)["segments"]
        ctx.sync()
        gpu_base = gpu_baseline(scene_path, w, h, spp, depth, seg_full, seg_d5)
        run.film.reset()
    sampler = ClockSampler(local_rank)
    sampler.start()
    m = run.measure(args.steps, args.warmup)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    clocks = sampler.summary()
    st_t = run.step(False, stage_timing=True)          # dominant-kernel time of one more frame (per-launch events; outside the timed regions)
    check = run.check_against_unsharded() if world > 1 and not args.no_check else None
    roof = roofline_of(run, m, st_t, args, clocks, sm_count) if pipeline in (pkg.PIPELINE_PERSIST, pkg.PIPELINE_FLAT, pkg.PIPELINE_DEFAULT) else None
    run.close()

    # ---- secondary records (outside the headline timing): the other configurations the targets are quoted on ----
    extra = {}
    if not args.no_extra and args.workload == "c2" and args.spp == 0:
        def secondary(workload, rng, spp_x, steps, note):
            _, ww, hh, spp_full, dd = WORKLOADS[workload]
            r = Run(pkg, ctx, comm, dist, torch, rank, world, workload, spp_x or spp_full, dd, rng, pkg.PIPELINE_DEFAULT, args)
            mm = r.measure(steps, 1, e2e=False)
            rec = {"workload": f"{workload}: {WORKLOADS[workload][0]} {ww}x{hh}, {spp_x or spp_full} spp, depth {dd}", "rng": rng, "value": mm["value"], "unit": METRIC,
                   "ms_per_step": mm["ms_per_step"], "steps": steps, "warmup": 1, "n_gpus": world, "note": note,
                   "sharding": {0: "none", 1: "tiles", 2: "sample ranges"}[r.shard_mode]}
            if world > 1:
                rec["multi_rank_check"] = r.check_against_unsharded()
            sst = r.scene.stats()
            if sst["node_bytes"] + sst["triangle_bytes"] > 126e6:
                stt = r.step(False, stage_timing=True)
                rec["roofline"] = roofline_of(r, mm, stt, args, clocks, sm_count)
                rec["scene"] = {k: sst[k] for k in ("n_triangles", "n_nodes", "build_ms", "sah_cost")}
            r.close()
            return rec
        extra["philox"] = secondary("c2", "philox", 0, 3, "counter RNG; sharded by sample range when n_gpus > 1")
        extra["c4"] = secondary("c4", "compat", 1024, 3, "the 4K mirror / emissive Cornell scene of the >= 7x scaling target, at 1024 of its 4096 spp")
        if world == 1:
            extra["c3"] = secondary("c3", "compat", 0, 3, "10.5 M triangles: the traversal-roofline target (>= 0.60)")

    if rank == 0:
        out = {
            "metric": METRIC, "value": m["value"], "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload_desc, "rng": args.rng, "pipeline": args.pipeline,
                       "sharding": {0: "none", 1: "tiles", 2: "sample ranges"}[run.shard_mode],
                       "film_combine": None if world == 1 else ("ptx_film_combine (library): pack owned tiles + ncclAllGather + unpack" if run.shard_mode == pkg.SHARD_TILES else "ptx_film_combine (library): in-place ncclAllReduce"),
                       "l2": "flushed before every step: 256 MB memset on the render stream, inside the timed region (film + RNG state are %.0f MB)" % (w * h * 40 / 1e6 / world),
                       "segments_per_step": m["seg"] / args.steps},
            "clocks": clocks,
            "e2e": {"value": m["e2e_value"], "unit": METRIC, "h2d_bytes_per_step": ctypes.sizeof(pkg.Camera), "d2h_bytes_per_step": w * h * 3,
                    "ms_per_step": m["ms_e2e"] / args.steps},
            "gpu_launches": m["launches"],
            "culling": {"what": "whole-pixel culling in ptx_render (on by default): pixels none of whose camera rays can reach the scene's bounding box are not traced; "
                                "their XORWOW streams are advanced by 2 * spp draws and their film sums left alone, so film and RNG state equal the traced frame's bit for bit "
                                "(tests: test_whole_pixel_culling_keeps_film_and_rng_state, the full-size comparisons with the reference's CUDA kernels). "
                                "`value` and `e2e.value` count the WORKLOAD's path segments (what the reference traces for this frame: its kernels trace every one); "
                                "value_traced_only counts only the segments this build traced; the roofline is computed from the traced ones",
                        "segments_per_step": m["seg_per_step"], "culled_paths_per_step": m["culled_per_step"],
                        "traced_segments_per_step": m["seg_per_step"] - m["culled_per_step"],
                        "value_traced_only": m["value_traced"], "e2e_value_traced_only": m["e2e_value_traced"]},
            "roofline": roof,
            "scene": {k: scene_stats[k] for k in ("n_triangles", "n_nodes", "node_bytes", "triangle_bytes", "build_ms", "builder", "max_depth", "sah_cost", "flat_units")},
        }
        if check is not None:
            out["multi_rank_check"] = check
        if extra:
            out["extra"] = extra
        if gpu_base is not None:
            out["gpu_baseline"] = gpu_base
            if gpu_base.get("value"):
                out["gpu_baseline"]["speedup_device"] = m["value"] / gpu_base["value"] / world
                out["gpu_baseline"]["note"] = "speedup_device = value / n_gpus / gpu_baseline.value (per-GPU throughput of this build over the reference CUDA build on one B200)"
        if world == 1 and not args.no_cpu_baseline and generated:
            out["cpu_baseline"] = {"value": None, "unit": METRIC, "cores": 0, "kind": "reference",
                                   "sample": "not runnable: the reference keeps 4 364-byte primitives (45.8 GB for this scene) and builds its BVH on one host thread (BASELINE.md)"}
        elif world == 1 and not args.no_cpu_baseline:
            try:
                cb_spp = 2
                spp_ratio = segments_per_path_estimate(scene_path, w, h, depth)
                j = ref_host_run(scene_path, w, h, cb_spp, depth, 1)
                # the same build on ONE thread as well -- on a 1/4-scale frame (render_init = curand_init with a per-pixel
                # subsequence jump costs 12 us per pixel on the host: 25 s for 1080p on one thread), 8 spp
                w1, h1, spp1 = max(w // 4, 1), max(h // 4, 1), 8
                j1 = ref_host_run(scene_path, w1, h1, spp1, depth, 1, threads=1)
                out["cpu_baseline"] = {"value": w * h * cb_spp * spp_ratio / (j["ms_per_frame"] * 1e-3) * 1e-6, "unit": METRIC, "cores": j["threads"],
                                       "single_thread_value": w1 * h1 * spp1 * spp_ratio / (j1["ms_per_frame"] * 1e-3) * 1e-6,
                                       "single_thread_sample": f"{w1}x{h1}, {spp1} spp, one thread",
                                       "kind": "reference", "sample": f"{w}x{h}, {cb_spp} of {spp} spp, depth {depth}, oracle/_ref/ref_host (reference headers, host build, OpenMP rows)",
                                       "segments_per_path_source": "C oracle on a 1/8-scale frame at 4 spp (same streams), not counted on the timed frame"}
            except Exception as e:       # noqa: BLE001
                out["cpu_baseline"] = {"value": None, "error": str(e)[:200]}
        emit(out)
    if comm is not None:
        comm.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    if check is not None and not check["ok"]:
        raise SystemExit("multi-rank film check failed: " + json.dumps(check))

_REAL_STDOUT = None


def emit(obj):
    """The ONE JSON line of the contract, written to the process's real stdout (see main())."""
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(obj) + "\n").encode())


def compat_chunk_count(spp, override, pixels_per_lane):
    """Work items per pixel of the compat-RNG persistent kernel (the rule of csrc/api.cu ptx_render: U equal parts, the last one
    halved down to spp / 32; U = 2 with >= 8 pixels per resident lane or spp < 512, else 8).  Only used for the algorithmic-bytes estimate."""
    if spp <= 0:
        return 1
    if override > 0:
        return min(override, spp.bit_length())
    u = -override if override < 0 else (2 if (pixels_per_lane >= 8.0 or spp < 512) else 8)
    u = max(1, min(u, spp, 32))
    n, at = max(u, 1), spp * (u - 1) // u
    left, floor_sz = spp - at, max(1, spp // 32)
    while left > floor_sz and left // 2 >= 1:
        at += left - left // 2
        left = spp - at
        n += 1
    return n


def main():
    # Native libraries print banners on fd 1 (NCCL: "NCCL version ..." at the first communicator).  The contract is one JSON
    # line on stdout, so fd 1 is pointed at stderr for the whole run and the line goes to a saved copy of the real stdout.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rng", default="compat", choices=["compat", "philox"])
    ap.add_argument("--pipeline", default="auto", choices=["auto", "flat", "persist", "loop", "fused", "wavefront", "cta"],
                    help="auto = the library default: the flat unit-list kernel for tiny scenes, else the persistent BVH8 kernel")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override spp (reduced runs are NOT the headline)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="N > 1: skip the unsharded re-render that checks the combined film (a whole frame per rank: expensive on c5)")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary records (extra.philox / extra.c4 / extra.c3) of the default C2 run")
    ap.add_argument("--min-blocks", type=int, default=0, help="experiment: occupancy variant of the path-loop kernels (0 = 3x256 threads/SM, 4 = 4x256, 7 = 7x128)")
    ap.add_argument("--chunks", type=int, default=0, help="experiment: work items per pixel of the persistent kernel, compat RNG (0 = default 6, 1 = none)")
    ap.add_argument("--shade-batch", type=int, default=0, help="experiment: waiting lanes per warp that trigger a shading round (0 = default 12, 1 = at once)")
    ap.add_argument("--no-cull", action="store_true", help="experiment: trace every pixel (PTX_RENDER_NO_BOUNDS_CULL)")
    ap.add_argument("--no-cost-order", action="store_true", help="experiment: hand out the work items in film order instead of longest-expected first")
    ap.add_argument("--max-leaf", type=int, default=0, help="experiment: triangles per BVH8 leaf slot (1..3; 0 = library default)")
    args = ap.parse_args()
    scene, w, h, spp, depth = WORKLOADS[args.workload]
    if args.spp > 0:
        spp = args.spp
    args.workload_desc = f"{args.workload}: scenes/{scene} {w}x{h}, {spp} spp, max depth {depth}, reference Russian roulette (depth>2, cap 0.95), SAMPLING_BSDF"
    scene_path = os.path.join(ROOT, "scenes", scene)
    if args.impl == "reference" and scene.startswith("generated:"):
        if int(os.environ.get("RANK", "0")) == 0:
            emit({"impl": "reference", "unavailable": "reference cannot hold this scene (4 364 B per primitive, single-thread BVH build); see BASELINE.md"})
        return
    if args.impl == "reference":
        run_reference(args, scene_path, w, h, spp, depth)
    else:
        run_ours(args, w, h, spp, depth)


if __name__ == "__main__":
    main()
