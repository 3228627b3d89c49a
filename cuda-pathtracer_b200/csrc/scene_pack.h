// scene_pack.h -- host-side packing of reference-shaped primitives into the device layout inputs.
#pragma once
#include <vector>

#include "../../include/ptx.h"
#include "bvh_host.h"

namespace ptx {

struct PackedScene {
    std::vector<TriIn> tris;            // quads contribute two, sharing prim_id
    std::vector<f32x4> prim_shade;      // { n.xyz, material index }
    std::vector<f32x4> materials;       // 2 per material: { Kd.rgb, type }, { Ke.rgb, 0 }
    std::vector<uint32_t> quad_bits;    // bit p = primitive p is a Quad; empty when the scene has none
};
// the bitset for a primitive list (also used for lists that were preprocessed on the GPU)
void make_quad_bits(const ptx_primitive* prims, size_t n, std::vector<uint32_t>& out);

// primitives in the reference's numbering -> triangle soup + deduplicated material table
void pack_primitives(const ptx_primitive* prims, size_t n, PackedScene& out);
// triangle soup + explicit material table (ptx_scene_from_triangles)
void pack_triangles(const float* v, const float* normals, const uint32_t* mat_id, const ptx_material* mats, size_t n_tri,
                    size_t n_mat, PackedScene& out);

}  // namespace ptx
