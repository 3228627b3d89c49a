// geometry_gpu.h -- geometry preprocessing on the device, feeding the BVH builder without a host round trip:
// convertQuadsToTriangles (application_state.h:323-365) and subdivide_primitives (form_factors.h:475-574) as CUDA
// kernels, then the split into the builder's triangle soup + per-primitive shading records.  Same arithmetic, same
// output order as the host versions in loader.cpp / scene_pack.cpp (tests compare them bit for bit).
#pragma once
#include <cuda_runtime.h>

#include <string>

#include "../../include/ptx.h"
#include "bvh_common.h"

namespace ptx {

struct DevGeometry {
    ptx_primitive* d_prims = nullptr;   // final primitives, reference numbering
    uint32_t* d_mat = nullptr;          // material index per final primitive
    size_t n_prims = 0;
    TriIn* d_tris = nullptr;            // builder input: quads contribute two, sharing prim_id
    size_t n_tris = 0;
    f32x4* d_prim_shade = nullptr;      // { n.xyz, material index }
    float ms = 0.f;                     // device time of the whole preprocessing
    void release();
};

// h_base: n0 primitives as parsed; h_base_mat: their material indices.  Everything else happens on the device.
cudaError_t preprocess_geometry_gpu(cudaStream_t st, const ptx_primitive* h_base, const uint32_t* h_base_mat, size_t n0,
                                    bool convert_quads, int levels, DevGeometry& out, std::string& err);

}  // namespace ptx
