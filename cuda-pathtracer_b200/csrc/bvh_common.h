// bvh_common.h -- per-node pieces of the BVH8 construction shared by the host builder (bvh_host.cpp)
// and the device builder (bvh_gpu.cu): choosing the <= 8 children of a wide node from a BVH2 subtree,
// octant-aware slot assignment, 80-byte node compression and the 48-byte triangle record.
// Written __host__ __device__ so that the CPU tests of the host builder exercise the very code the
// GPU collapse kernels run.
#pragma once
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

struct WideChild {
    aabb box;
    uint32_t is_inner;       // 1 inner, 0 leaf
    uint32_t tri_count;      // leaf only, 1..3
    uint32_t tri_offset;     // leaf only: offset from the node's tri_base
};

// Padding of every primitive's box before the build.  The reference pads its BVH2 boxes by 1e-6 (rendering/bvh.h:110-150,
// restated at oracle/pt_oracle.c:181): a ray aimed exactly at a vertex or an edge that lies on a box face is then still inside
// the box when the exact triangle test -- which alone decides a hit -- accepts it within rounding.  The relative term
// (4 ulp of the coordinate) keeps that true for scenes far from the origin, where 1e-6 is below one ulp.
PTX_HD float box_pad(float lo, float hi) { return fmaxf(1e-6f, 4.76837158e-7f * fmaxf(fabsf(lo), fabsf(hi))); }

constexpr float kTravCost = 1.0f;   // cost of visiting one wide-node child box relative to one triangle test

PTX_HD void pack_triangle(const TriIn& t, f32x4 out[3]) {
    vec3 v0 = ld3(t.v0), e1 = ld3(t.v1) - v0, e2 = ld3(t.v2) - v0;
    out[0] = f32x4{v0.x, v0.y, v0.z, u2f(t.prim_id)};
    out[1] = f32x4{e1.x, e1.y, e1.z, 0.f};
    out[2] = f32x4{e2.x, e2.y, e2.z, 0.f};
}

// Greedy surface-area collapse: start from the node's two children and keep opening the inner child
// with the largest area until there are 8 children or only leaves.  Returns the child count.
PTX_HD int gather_wide_children(const Bvh2Node* nodes, uint32_t root, uint32_t kids[8]) {
    const Bvh2Node& nd = nodes[root];
    int nk = 0;
    if (nd.count > 0) kids[nk++] = root;                           // degenerate: the subtree root itself is a leaf
    else { kids[nk++] = nd.left; kids[nk++] = nd.right; }
    while (nk < 8) {
        int pick = -1; float best = -1.f;
        for (int i = 0; i < nk; i++) {
            const Bvh2Node& c = nodes[kids[i]];
            if (c.count > 0) continue;
            float a = aabb_half_area(c.box);
            if (a > best) { best = a; pick = i; }
        }
        if (pick < 0) break;
        uint32_t open = kids[pick];
        kids[pick] = nodes[open].left;
        kids[nk++] = nodes[open].right;
    }
    return nk;
}

// Slot assignment: slot s = corner (x: bit 2, y: bit 1, z: bit 0) of the node the child sits in, so that
// visiting slots by descending (s ^ octinv) is approximately front to back.  Greedy maximum of
// dot(child centre - node centre, corner direction).  child_in_slot[s] = index into kids or -1.
PTX_HD void assign_slots(const Bvh2Node* nodes, const aabb& node_box, const uint32_t* kids, int nk, int child_in_slot[8]) {
    vec3 nc = mk3(0.5f * (node_box.lo.x + node_box.hi.x), 0.5f * (node_box.lo.y + node_box.hi.y), 0.5f * (node_box.lo.z + node_box.hi.z));
    float cost[8][8];
    for (int i = 0; i < nk; i++) {
        const aabb& cb = nodes[kids[i]].box;
        vec3 dlt = mk3(0.5f * (cb.lo.x + cb.hi.x), 0.5f * (cb.lo.y + cb.hi.y), 0.5f * (cb.lo.z + cb.hi.z)) - nc;
        for (int s = 0; s < 8; s++)
            cost[i][s] = ((s & 4) ? dlt.x : -dlt.x) + ((s & 2) ? dlt.y : -dlt.y) + ((s & 1) ? dlt.z : -dlt.z);
    }
    int slot_of[8];
    for (int i = 0; i < 8; i++) { slot_of[i] = -1; child_in_slot[i] = -1; }
    for (int round = 0; round < nk; round++) {
        float best = -FLT_MAX; int bi = -1, bs = -1;
        for (int i = 0; i < nk; i++) {
            if (slot_of[i] >= 0) continue;
            for (int s = 0; s < 8; s++) {
                if (child_in_slot[s] >= 0) continue;
                if (cost[i][s] > best) { best = cost[i][s]; bi = i; bs = s; }
            }
        }
        slot_of[bi] = bs; child_in_slot[bs] = bi;
    }
}

// Quantisation frame + child boxes.  Exponent e (biased byte) is the smallest with
// ceil((hi - lo) / 2^(e-127)) <= 255; lo is floored and hi ceiled so decoded boxes contain the originals.
PTX_HD void encode_node(const aabb& nb, const WideChild* slots /*8, box.lo.x > box.hi.x => empty*/, uint32_t child_base,
                        uint32_t tri_base, u32x4 out[5]) {
    float p[3] = {nb.lo.x, nb.lo.y, nb.lo.z};
    float ext[3] = {nb.hi.x - nb.lo.x, nb.hi.y - nb.lo.y, nb.hi.z - nb.lo.z};
    uint32_t e[3]; float inv_scale[3];
    for (int a = 0; a < 3; a++) {
        uint32_t eb = 0;
        if (ext[a] > 0.f) {
            int ex; frexpf(ext[a] / 255.0f, &ex);             // ext/255 = m * 2^ex, m in [0.5,1)  => 2^ex >= ext/255
            int be = ex + 127;
            if (be < 1) be = 1;
            if (be > 254) be = 254;
            eb = (uint32_t)be;
            while (eb < 254 && ceilf(ext[a] / u2f(eb << 23)) > 255.0f) eb++;     // guard against rounding in the division above
        }
        e[a] = eb;
        inv_scale[a] = eb ? 1.0f / u2f(eb << 23) : 0.f;
    }
    uint32_t meta[8], q[6][8];
    uint32_t imask = 0;
    for (int s = 0; s < 8; s++) {
        const WideChild& c = slots[s];
        bool empty = c.box.lo.x > c.box.hi.x;
        if (empty) { meta[s] = 0; for (int k = 0; k < 6; k++) q[k][s] = 0; continue; }
        if (c.is_inner) { meta[s] = (1u << 5) | (24u + (uint32_t)s); imask |= 1u << s; }
        else { uint32_t unary = c.tri_count == 1 ? 1u : (c.tri_count == 2 ? 3u : 7u); meta[s] = (unary << 5) | (c.tri_offset & 31u); }
        float lo[3] = {c.box.lo.x, c.box.lo.y, c.box.lo.z}, hi[3] = {c.box.hi.x, c.box.hi.y, c.box.hi.z};
        for (int a = 0; a < 3; a++) {
            float scale = e[a] ? u2f(e[a] << 23) : 0.f;
            float ql = floorf((lo[a] - p[a]) * inv_scale[a]);
            float qh = ceilf((hi[a] - p[a]) * inv_scale[a]);
            ql = ql < 0.f ? 0.f : (ql > 255.f ? 255.f : ql);
            qh = qh < 0.f ? 0.f : (qh > 255.f ? 255.f : qh);
            // make the decoded box conservative under float rounding of p + q*scale
            while (ql > 0.f && p[a] + ql * scale > lo[a]) ql -= 1.f;
            while (qh < 255.f && p[a] + qh * scale < hi[a]) qh += 1.f;
            q[a][s] = (uint32_t)ql; q[3 + a][s] = (uint32_t)qh;
        }
    }
#define PTX_PACK4(b, o) ((b)[(o)] | ((b)[(o) + 1] << 8) | ((b)[(o) + 2] << 16) | ((b)[(o) + 3] << 24))
    out[0] = u32x4{f2u(p[0]), f2u(p[1]), f2u(p[2]), e[0] | (e[1] << 8) | (e[2] << 16) | (imask << 24)};
    out[1] = u32x4{child_base, tri_base, PTX_PACK4(meta, 0), PTX_PACK4(meta, 4)};
    out[2] = u32x4{PTX_PACK4(q[0], 0), PTX_PACK4(q[0], 4), PTX_PACK4(q[1], 0), PTX_PACK4(q[1], 4)};
    out[3] = u32x4{PTX_PACK4(q[2], 0), PTX_PACK4(q[2], 4), PTX_PACK4(q[3], 0), PTX_PACK4(q[3], 4)};
    out[4] = u32x4{PTX_PACK4(q[4], 0), PTX_PACK4(q[4], 4), PTX_PACK4(q[5], 0), PTX_PACK4(q[5], 4)};
#undef PTX_PACK4
}

// One wide node: fills `slots` (leaf children get consecutive triangle offsets in slot order), the inner
// children in slot order (`inner_kids`), and the leaf children's BVH2 nodes in slot order (`leaf_kids`).
struct WideNodePlan {
    WideChild slots[8];
    uint32_t inner_kids[8];
    uint32_t leaf_kids[8];
    int n_inner, n_leaf;
    uint32_t n_tris;
    float leaf_area_cost;     // sum over leaf children of half_area * count (not normalised)
};

PTX_HD void plan_wide_node(const Bvh2Node* nodes, uint32_t root, WideNodePlan& plan) {
    uint32_t kids[8];
    int nk = gather_wide_children(nodes, root, kids);
    int child_in_slot[8];
    assign_slots(nodes, nodes[root].box, kids, nk, child_in_slot);
    plan.n_inner = 0; plan.n_leaf = 0; plan.n_tris = 0; plan.leaf_area_cost = 0.f;
    for (int s = 0; s < 8; s++) {
        WideChild& w = plan.slots[s];
        w.box = aabb_empty(); w.is_inner = 0; w.tri_count = 0; w.tri_offset = 0;
        int ci = child_in_slot[s];
        if (ci < 0) continue;
        const Bvh2Node& c = nodes[kids[ci]];
        w.box = c.box;
        if (c.count > 0) {
            w.tri_count = c.count; w.tri_offset = plan.n_tris;
            plan.n_tris += c.count;
            plan.leaf_kids[plan.n_leaf++] = kids[ci];
            plan.leaf_area_cost += aabb_half_area(c.box) * (float)c.count;
        } else {
            w.is_inner = 1;
            plan.inner_kids[plan.n_inner++] = kids[ci];
        }
    }
}

}  // namespace ptx
