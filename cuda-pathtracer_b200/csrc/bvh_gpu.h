// bvh_gpu.h -- device BVH builder (bvh_gpu.cu): binned-SAH BVH2 built breadth-first on the GPU, collapsed and
// compressed to the BVH8 layout of layout.h without leaving HBM.
#pragma once
#include <cuda_runtime.h>

#include <string>

#include "bvh_common.h"

namespace ptx {

struct GpuBvhResult {
    u32x4* d_nodes = nullptr;      // n_nodes * 5, cudaMalloc'ed; owned by the caller
    f32x4* d_tris = nullptr;       // n_tris * 3
    uint32_t n_nodes = 0, n_tris = 0;
    uint32_t n_bvh2_nodes = 0;
    int bvh2_levels = 0;
    int max_depth = 0;             // BVH8 depth
    float sah_cost = 0.f;
    aabb bounds;
    float stage_ms[4] = {0.f, 0.f, 0.f, 0.f};   // upload + boxes + Morton sort | binned-SAH levels | leaf order | BVH8 collapse
};

// tris: n input triangles, on the host (uploaded here) or already on the device (tris_on_device; read in place, e.g. the
// output of preprocess_geometry_gpu).  Returns cudaSuccess or the failing CUDA error; `err` gets the text.
// A tree deeper than `max_bvh8_depth` (the traversal stack size) is reported as cudaErrorInvalidValue.
cudaError_t build_bvh8_gpu(cudaStream_t stream, const TriIn* tris, size_t n, int max_leaf, int max_bvh8_depth, GpuBvhResult& out,
                           std::string& err, bool tris_on_device = false);

}  // namespace ptx
