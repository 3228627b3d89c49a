"""Build recipe for libptx.so (the C-ABI library) -- explicit nvcc for sm_100a (one object per source, then a link), built in-tree.

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
HOST_SOURCES = ["bvh_host.cpp", "flat_host.cpp", "image_io.cpp", "loader.cpp", "pbrt_loader.cpp", "rng.cpp", "scene_pack.cpp"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-ffp-contract=off", "-diag-suppress", "177",
]


def _newer_than_lib(paths):
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in paths)


def _compile_one(nvcc, src, obj, verbose):
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas=-v"] if verbose else []) + ["-c", "-o", obj, src]
    r = subprocess.run(cmd, capture_output=True, text=True)
    return src, r


def build_lib(force=False, verbose=False):
    """One object per source (compiled in parallel, rebuilt only when the source or any header is newer), then one link."""
    from concurrent.futures import ThreadPoolExecutor
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    srcs = [os.path.join(CSRC, s) for s in CUDA_SOURCES + HOST_SOURCES if os.path.exists(os.path.join(CSRC, s))]
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "ptx.h"))
    if not force and not _newer_than_lib(srcs + headers):
        return LIB
    if not os.path.exists(nvcc):
        if os.path.exists(LIB):
            return LIB          # GPU box: use the prebuilt library that travelled with the snapshot
        raise RuntimeError("nvcc not found and no prebuilt libptx.so")
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    hdr_time = max(os.path.getmtime(h) for h in headers)
    todo, objs = [], []
    for src in srcs:
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_time):
            todo.append((src, obj))
    with ThreadPoolExecutor(max_workers=max(1, min(len(todo), os.cpu_count() or 4))) as ex:
        for src, r in ex.map(lambda so: _compile_one(nvcc, so[0], so[1], verbose), todo):
            if r.returncode != 0:
                sys.stderr.write(r.stdout + r.stderr)
                raise RuntimeError(f"nvcc failed compiling {os.path.basename(src)}")
            if verbose:
                print(r.stderr)
    cmd = [nvcc] + NVCC_FLAGS + ["-shared", "-o", LIB] + objs
    if verbose:
        print(" ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed linking libptx.so")
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
