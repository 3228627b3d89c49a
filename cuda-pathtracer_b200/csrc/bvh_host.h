// bvh_host.h -- host-side binned-SAH builder + BVH8 collapse/compression.
//
// This is the CROSS-CHECK builder (ptx_build_opts.builder = 1): it emits exactly the node / triangle
// layout of layout.h, so the GPU builder's output can be validated against it (same hits) and scenes
// can be brought up before the device builder runs.  It replaces the reference's BVHBuilder
// (rendering/bvh.h:76-219: recursive, longest-centroid-axis midpoint split, leaf <= 4).
#pragma once
#include <cstdint>
#include <vector>

#include "bvh_common.h"

namespace ptx {

struct Bvh2 {
    std::vector<Bvh2Node> nodes;      // nodes[0] = root
    std::vector<uint32_t> order;      // triangle indices in leaf order
};

struct Bvh8Result {
    std::vector<u32x4> nodes;         // 5 per node
    std::vector<f32x4> tris;          // 3 per triangle, leaf order
    std::vector<uint32_t> tri_src;    // tri_src[slot] = index of the input triangle
    float sah_cost = 0.f;
    int max_depth = 0;
    aabb bounds;
};

void build_bvh2_sah(const TriIn* tris, size_t n, int bins, int max_leaf, Bvh2& out);
void collapse_bvh8(const Bvh2& b2, const TriIn* tris, size_t n, Bvh8Result& out);

}  // namespace ptx
