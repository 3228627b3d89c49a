// bvh_host.h -- host-side binned-SAH builder + BVH8 collapse/compression.
//
// This is the CROSS-CHECK builder (ptx_build_opts.builder = 1): it emits exactly the node / triangle
// layout of layout.h, so the GPU builder's output can be validated against it (same hits) and scenes
// can be brought up before the device builder runs.  It replaces the reference's BVHBuilder
// (rendering/bvh.h:76-219: recursive, longest-centroid-axis midpoint split, leaf <= 4).
#pragma once
#include <cstdint>
#include <vector>

#include "layout.h"

namespace ptx {

struct TriIn {          // one input triangle: vertices + the reference primitive it belongs to
    float v0[3], v1[3], v2[3];
    uint32_t prim_id;
};

struct Bvh2Node {
    aabb box;
    uint32_t left, right;    // children (inner) ; leaf: left = first index into order[], right = 0
    uint32_t count;          // > 0 => leaf
};

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

// shared with the device builder's tests: quantise one wide node
struct WideChild {
    aabb box;
    uint32_t is_inner;       // 1 inner, 0 leaf
    uint32_t tri_count;      // leaf only, 1..3
    uint32_t tri_offset;     // leaf only: offset from the node's tri_base
};
void encode_node(const aabb& node_box, const WideChild* slots /*8, box.lo.x > box.hi.x => empty*/, uint32_t child_base,
                 uint32_t tri_base, u32x4 out[5]);
void pack_triangle(const TriIn& t, f32x4 out[3]);

}  // namespace ptx
