"""Regenerates tests/golden/*.  Run in the build container (needs /root/reference, via oracle/_ref/ref_host).

Every vector here is produced by the REFERENCE'S OWN CODE compiled for the host (oracle/_ref/ref_host:
the unmodified headers under /root/reference/include driven by oracle/ref_harness.cu), except
converged_32.npz which is produced by the plain-C oracle after that oracle has been pinned bit-exactly
to ref_host by the other vectors (ref_host only returns per-frame means; the oracle also returns the
second moment needed for a Monte-Carlo-noise test).

    python tests/golden/make_golden.py
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "ref_host")
SCENES = os.path.join(ROOT, "scenes")
sys.path.insert(0, os.path.join(ROOT, "tests"))


def run(*args):
    r = subprocess.run([REF] + [str(a) for a in args], capture_output=True, text=True, check=True)
    return json.loads(r.stdout.strip().splitlines()[-1])


def main():
    tmp = tempfile.mkdtemp()
    out = {}
    # 1. XORWOW known answers: curand_init(seed, subsequence, 0) state + first 8 curand_uniform draws
    kats = []
    for seed, sub in [(2030, 7), (2023, 0), (2024, 1), (2023 + 262143, 262143), (2023 + 2073599, 2073599), (2023 + 8294399, 8294399),
                      (12345678901, 1099511640121), (0, 3), (2**40 + 5, 65536)]:
        j = run("rng", seed, sub, 8)
        kats.append({"seed": seed, "subsequence": sub, "v": j["v"], "d": j["d"], "uniform": j["uniform"]})
    with open(os.path.join(HERE, "xorwow_kat.json"), "w") as f:
        json.dump(kats, f, indent=1)

    arrays = {}
    for name in ("cbox", "cbox_quads"):
        path = os.path.join(SCENES, name + ".obj")
        for cq in (0, 1):
            if name == "cbox" and cq:
                continue
            tag = f"{name}_cq{cq}"
            # 2. loader output (primitive records) and the reference BVH2 (node count is a loader KAT too)
            j = run("dump", path, os.path.join(tmp, tag), cq)
            arrays[tag + "_prims"] = np.fromfile(os.path.join(tmp, tag + ".prims.bin"), np.float32).reshape(-1, 22)
            arrays[tag + "_nodes"] = np.fromfile(os.path.join(tmp, tag + ".nodes.bin"), np.float32).reshape(-1, 9)
            arrays[tag + "_indices"] = np.fromfile(os.path.join(tmp, tag + ".indices.i32"), np.int32)
            # 3. primary hit ids at pixel centres
            for w, h in ((64, 64), (160, 90)):
                run("ids", path, w, h, os.path.join(tmp, tag), cq)
                arrays[f"{tag}_ids_{w}x{h}"] = np.fromfile(os.path.join(tmp, tag + ".ids.i32"), np.int32).reshape(h, w).astype(np.int8)
                arrays[f"{tag}_t_{w}x{h}"] = np.fromfile(os.path.join(tmp, tag + ".t.f32"), np.float32).reshape(h, w)
            # 4. fixed-seed radiance / RGB8
            for (w, h, spp, depth) in ((48, 48, 8, 8), (64, 36, 4, 16), (32, 32, 16, 5)):
                j = run("render", path, w, h, spp, depth, 1, os.path.join(tmp, tag), cq, 0)
                key = f"{tag}_{w}x{h}_s{spp}_d{depth}"
                arrays[key + "_linear"] = np.fromfile(os.path.join(tmp, tag + ".linear.f32"), np.float32).reshape(h, w, 3)
                arrays[key + "_rgb8"] = np.fromfile(os.path.join(tmp, tag + ".rgb8.u8"), np.uint8).reshape(h, w, 3)
            # two frames: the stream persists across frames
            j = run("render", path, 32, 32, 4, 5, 2, os.path.join(tmp, tag), cq, 0)
            arrays[f"{tag}_32x32_s4_d5_frame2_linear"] = np.fromfile(os.path.join(tmp, tag + ".linear.f32"), np.float32).reshape(32, 32, 3)
        # 5. random rays: BVH traversal and linear scan of the reference
        rng = np.random.default_rng(11)
        n = 3000
        o = rng.uniform([-2.5, 0.2, -5.0], [2.0, 5.0, 6.0], (n, 3))
        d = rng.normal(size=(n, 3))
        rays = np.concatenate([o, d], 1).astype(np.float32)
        rays.tofile(os.path.join(tmp, "rays.bin"))
        j = run("rays", path, os.path.join(tmp, "rays.bin"), n, os.path.join(tmp, name), 0)
        arrays[name + "_rays"] = rays
        arrays[name + "_rays_bvh_ids"] = np.fromfile(os.path.join(tmp, name + ".bvh.i32"), np.int32).astype(np.int8)
        arrays[name + "_rays_bvh_t"] = np.fromfile(os.path.join(tmp, name + ".bvh_t.f32"), np.float32)
        arrays[name + "_rays_lin_ids"] = np.fromfile(os.path.join(tmp, name + ".lin.i32"), np.int32).astype(np.int8)
        arrays[name + "_rays_lin_t"] = np.fromfile(os.path.join(tmp, name + ".lin_t.f32"), np.float32)
        out[name + "_bvh_vs_linear_mismatch"] = j["bvh_vs_linear_mismatch"]
    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **arrays)

    # 6. converged image: 65 536 spp, 32x32, depth 8 (oracle, pinned by the vectors above)
    import oracle_binding
    orc = oracle_binding.Oracle()
    conv = {}
    for name in ("cbox", "cbox_quads"):
        s = orc.load_obj(os.path.join(SCENES, name + ".obj"))
        sm, sq = orc.render_moments(s, 32, 32, 65536, 8)
        conv[name + "_mean"] = (sm / 65536).astype(np.float32)
        conv[name + "_var"] = np.maximum(sq / 65536 - (sm / 65536) ** 2, 0).astype(np.float32)   # per-sample variance
        orc.free(s)
    np.savez_compressed(os.path.join(HERE, "converged_32.npz"), **conv)
    with open(os.path.join(HERE, "meta.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("golden vectors written:", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
