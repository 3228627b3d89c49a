"""Build recipe for libptx.so (the C-ABI library) -- explicit nvcc for sm_100a, built in-tree.

`python cuda-pathtracer_b200/build.py` or `__graft_entry__.build()`.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libptx.so")

CUDA_SOURCES = ["api.cu", "kernels.cu", "bvh_gpu.cu", "geometry_gpu.cu"]
HOST_SOURCES = ["bvh_host.cpp", "image_io.cpp", "loader.cpp", "pbrt_loader.cpp", "rng.cpp", "scene_pack.cpp"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-ffp-contract=off", "-diag-suppress", "177",
]


def _newer_than_lib(paths):
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in paths)


def build_lib(force=False, verbose=False):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    srcs = [os.path.join(CSRC, s) for s in CUDA_SOURCES + HOST_SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    deps.append(os.path.join(HERE, "..", "include", "ptx.h"))
    if not force and not _newer_than_lib(deps):
        return LIB
    if not os.path.exists(nvcc):
        if os.path.exists(LIB):
            return LIB          # GPU box: use the prebuilt library that travelled with the snapshot
        raise RuntimeError("nvcc not found and no prebuilt libptx.so")
    cmd = [nvcc] + NVCC_FLAGS + ["-shared", "-o", LIB] + srcs
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building libptx.so")
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
