// bvh_gpu.cu -- device BVH builder: breadth-first binned-SAH BVH2, then level-synchronous collapse to
// compressed BVH8 nodes + 48-byte triangle records (layout.h), everything resident in HBM.
//
// Replaces the reference's host-side recursive BVHBuilder (rendering/bvh.h:76-219: longest-centroid-axis
// midpoint split, leaf <= 4, single thread) and the cudaMemcpy of its nodes / index permutation
// (rendering/ray_tracing_backend.h:81-129).
//
// Pipeline (all on the caller's stream):
//   1. k_prim_setup      triangle AABB + centroid, scene / centroid bounds (warp-reduced float atomics)
//   2. Morton sort       30-bit codes of the centroids, cub radix sort -> processing order (locality for the atomics)
//   3. per level, over ALL primitives, no data movement (a primitive only carries the id of its current node):
//        k_bin           16 bins on each axis of the node's centroid bounds: box + count per bin (atomics)
//        k_split         one thread per node of the level: SAH sweep, leaf decision, child boxes / counts
//        scan            exclusive sum of the split flags -> deterministic child indices
//        k_make_children write the two children
//        k_relabel       primitive -> child id; child centroid bounds and Morton-rank range (atomics)
//   4. leaf order        stable radix sort of (leaf id, rank) -> triangles of a leaf contiguous
//   5. collapse          per BVH8 level: plan (shared host/device code, bvh_common.h), scan, emit nodes + triangles
//
// Determinism: child indices and BVH8 node / triangle positions come from prefix sums, never from the order
// in which atomics land; float min/max atomics are order-independent.
#include <cub/cub.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <string>
#include <vector>

#include "bvh_gpu.h"

namespace ptx {

namespace {

constexpr int kThreads = 256;
constexpr int kBins = 16;

// order-preserving float <-> uint map, so that float min / max are integer atomics
__device__ __forceinline__ uint32_t f2ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

// Box whose fields grow by atomicMax only and whose EMPTY state is all zero bits (so cudaMemset clears it):
// hi[a] holds f2ord(hi), lo[a] holds ~f2ord(lo).
struct OrdBox {            // 24 B
    uint32_t lo[3], hi[3];
};
__device__ __forceinline__ void ordbox_atomic_grow(OrdBox* b, const float lo[3], const float hi[3]) {
    for (int a = 0; a < 3; a++) { atomicMax(&b->lo[a], ~f2ord(lo[a])); atomicMax(&b->hi[a], f2ord(hi[a])); }
}
__device__ __forceinline__ float ordbox_lo(const OrdBox& b, int a) { return ord2f(~b.lo[a]); }
__device__ __forceinline__ float ordbox_hi(const OrdBox& b, int a) { return ord2f(b.hi[a]); }

struct PrimBox {           // 32 B: AABB of one triangle (centroid = box centre, recomputed on the fly)
    float lo[3], hi[3];
    uint32_t tri;          // index of the input triangle
    uint32_t pad;
};

// build-time BVH2 node (device): Bvh2Node plus what the next level needs
struct BuildNode {
    OrdBox cbox;           // centroid bounds (filled by atomics while the parent is relabelled)
    uint32_t rank_lo, rank_hi;   // min / max Morton rank of its primitives (fallback split)
    int32_t axis;          // -1: no centroid extent (leaf or rank split); >= 0: binned; after k_split the chosen split axis
    float cmin[3], cscale[3];    // bin on axis a = (centroid[a] - cmin[a]) * cscale[a]; cscale[a] == 0: axis not binned
    int32_t split_bin;     // primitives with bin <= split_bin go left; -2 => fallback split by rank
    uint32_t rank_mid;
    uint32_t pcount;       // primitives in the node
    OrdBox fbox;           // full bounds by atomics: only for children of a fallback (rank) split
    uint32_t box_from_atomics;
};

struct SplitChildren {     // written by k_split for nodes that split
    aabb lbox, rbox;
    uint32_t lcount, rcount;
};

struct Bin {               // 28 B
    OrdBox box;
    uint32_t count;
};

__global__ void __launch_bounds__(kThreads) k_prim_setup(const TriIn* __restrict__ tris, uint32_t n, PrimBox* boxes, OrdBox* scene /*[0]=box,[1]=cbox*/) {
    uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX}, c[3] = {0, 0, 0};
    bool valid = i < n;
    if (valid) {
        const TriIn t = tris[i];
        for (int a = 0; a < 3; a++) {
            lo[a] = fminf(t.v0[a], fminf(t.v1[a], t.v2[a]));
            hi[a] = fmaxf(t.v0[a], fmaxf(t.v1[a], t.v2[a]));
            c[a] = 0.5f * (lo[a] + hi[a]);
            float pad = box_pad(lo[a], hi[a]);                 // bvh_common.h: the reference's 1e-6 box padding
            lo[a] -= pad; hi[a] += pad;
        }
        PrimBox pb;
        for (int a = 0; a < 3; a++) { pb.lo[a] = lo[a]; pb.hi[a] = hi[a]; }
        pb.tri = i; pb.pad = 0;
        boxes[i] = pb;
    }
    // warp reduction, then one set of atomics per warp
    float clo[3], chi[3];
    for (int a = 0; a < 3; a++) { clo[a] = valid ? c[a] : FLT_MAX; chi[a] = valid ? c[a] : -FLT_MAX; }
    for (int o = 16; o > 0; o >>= 1)
        for (int a = 0; a < 3; a++) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xFFFFFFFFu, lo[a], o)); hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xFFFFFFFFu, hi[a], o));
            clo[a] = fminf(clo[a], __shfl_xor_sync(0xFFFFFFFFu, clo[a], o)); chi[a] = fmaxf(chi[a], __shfl_xor_sync(0xFFFFFFFFu, chi[a], o));
        }
    if ((threadIdx.x & 31) == 0 && lo[0] <= hi[0]) { ordbox_atomic_grow(&scene[0], lo, hi); ordbox_atomic_grow(&scene[1], clo, chi); }
}

__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void __launch_bounds__(kThreads) k_morton(const PrimBox* __restrict__ boxes, uint32_t n, const OrdBox* __restrict__ scene, uint32_t* keys,
                                                     uint32_t* vals) {
    uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    const PrimBox pb = boxes[i];
    uint32_t code = 0;
    for (int a = 0; a < 3; a++) {
        float lo = ordbox_lo(scene[1], a), hi = ordbox_hi(scene[1], a);
        float c = 0.5f * (pb.lo[a] + pb.hi[a]);
        float x = hi > lo ? (c - lo) / (hi - lo) : 0.f;
        uint32_t q = (uint32_t)fminf(fmaxf(x * 1024.0f, 0.0f), 1023.0f);
        code |= expand_bits10(q) << (2 - a);
    }
    keys[i] = code; vals[i] = i;
}

__global__ void __launch_bounds__(kThreads) k_gather_boxes(const PrimBox* __restrict__ in, const uint32_t* __restrict__ order, uint32_t n, PrimBox* out) {
    uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    if (i < n) out[i] = in[order[i]];
}


__device__ __forceinline__ OrdBox ordbox_empty() {
    OrdBox b;
    for (int a = 0; a < 3; a++) { b.lo[a] = 0u; b.hi[a] = 0u; }
    return b;
}
__device__ __forceinline__ aabb ordbox_to_aabb(const OrdBox& o) {
    aabb b;
    b.lo = mk3(ordbox_lo(o, 0), ordbox_lo(o, 1), ordbox_lo(o, 2));
    b.hi = mk3(ordbox_hi(o, 0), ordbox_hi(o, 1), ordbox_hi(o, 2));
    return b;
}

__global__ void k_init_root(Bvh2Node* nodes, BuildNode* bn, const OrdBox* scene, uint32_t n) {
    nodes[0].box = ordbox_to_aabb(scene[0]);
    nodes[0].left = 0; nodes[0].right = 0; nodes[0].count = 0;
    BuildNode b;
    b.cbox = scene[1]; b.rank_lo = 0; b.rank_hi = n - 1; b.axis = -1; for (int a = 0; a < 3; a++) { b.cmin[a] = 0.f; b.cscale[a] = 0.f; } b.split_bin = -1; b.rank_mid = 0;
    b.pcount = n; b.fbox = ordbox_empty(); b.box_from_atomics = 0;
    bn[0] = b;
}

// first kernel of a level: the node's centroid bounds are complete now -> choose the binning axis
__global__ void __launch_bounds__(kThreads) k_prepare(Bvh2Node* nodes, BuildNode* bn, uint32_t lb, uint32_t le) {
    uint32_t node = lb + blockIdx.x * kThreads + threadIdx.x;
    if (node >= le) return;
    BuildNode b = bn[node];
    if (b.box_from_atomics) nodes[node].box = ordbox_to_aabb(b.fbox);
    int axis = -1;
    for (int a = 0; a < 3; a++) {
        b.cmin[a] = 0.f; b.cscale[a] = 0.f;
        if (b.pcount > 1) {
            float lo = ordbox_lo(b.cbox, a), hi = ordbox_hi(b.cbox, a);
            if (hi > lo) { b.cmin[a] = lo; b.cscale[a] = (float)kBins / (hi - lo); axis = a; }
        }
    }
    b.axis = axis;
    b.split_bin = -1;
    bn[node] = b;
}

__device__ __forceinline__ int bin_of(const PrimBox& pb, int axis, float cmin, float cscale) {
    float c = 0.5f * (pb.lo[axis] + pb.hi[axis]);
    int k = (int)((c - cmin) * cscale);
    return k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
}

// Warp-aggregated growth of an OrdBox: the lanes of `group` (same destination; from __match_any_sync) reduce their six order-preserving
// words with one REDUX each and ONLY the group's first lane issues the atomics.  The primitives are in Morton order, so the 32 lanes of
// a warp mostly share a node and fall into a handful of bins: on the upper levels of the tree -- where all 10 M primitives hammer a few
// hundred addresses -- this removes ~ 90 % of the atomics and nearly all of their contention.
__device__ __forceinline__ void ordbox_group_grow(OrdBox* dst, unsigned group, bool leader, const float lo[3], const float hi[3]) {
    uint32_t wlo[3], whi[3];
#pragma unroll
    for (int a = 0; a < 3; a++) { wlo[a] = __reduce_max_sync(group, ~f2ord(lo[a])); whi[a] = __reduce_max_sync(group, f2ord(hi[a])); }
    if (leader) {
#pragma unroll
        for (int a = 0; a < 3; a++) { atomicMax(&dst->lo[a], wlo[a]); atomicMax(&dst->hi[a], whi[a]); }
    }
}

__global__ void __launch_bounds__(kThreads) k_bin(const PrimBox* __restrict__ boxes, const uint32_t* __restrict__ prim_node, uint32_t n,
                                                  const BuildNode* __restrict__ bn, uint32_t lb, uint32_t le, Bin* bins) {
    uint32_t r = blockIdx.x * kThreads + threadIdx.x;
    bool active = r < n;
    uint32_t node = active ? prim_node[r] : 0u;
    active = active && node >= lb && node < le;
    const BuildNode* b = bn + node;
    active = active && b->axis >= 0;
    if (!__any_sync(0xFFFFFFFFu, active)) return;
    PrimBox pb;
    if (active) pb = boxes[r];
    const int lane = (int)(threadIdx.x & 31);
    for (int a = 0; a < 3; a++) {
        const float cs = active ? b->cscale[a] : 0.f;
        const bool on = active && cs > 0.f;
        const unsigned live = __ballot_sync(0xFFFFFFFFu, on);
        if (!on) continue;
        const uint32_t slot = ((node - lb) * 3u + (uint32_t)a) * (uint32_t)kBins + (uint32_t)bin_of(pb, a, b->cmin[a], cs);
        const unsigned group = __match_any_sync(live, slot);
        const bool leader = lane == __ffs(group) - 1;
        Bin* dst = bins + slot;
        ordbox_group_grow(&dst->box, group, leader, pb.lo, pb.hi);
        if (leader) atomicAdd(&dst->count, (uint32_t)__popc(group));
    }
}

__global__ void __launch_bounds__(kThreads) k_split(Bvh2Node* nodes, BuildNode* bn, uint32_t lb, uint32_t le, const Bin* __restrict__ bins,
                                                    int max_leaf, uint32_t* flags, SplitChildren* sc) {
    uint32_t node = lb + blockIdx.x * kThreads + threadIdx.x;
    if (node >= le) return;
    BuildNode b = bn[node];
    uint32_t cnt = b.pcount;
    uint32_t flag = 0;
    if (b.axis < 0) {
        if (cnt > (uint32_t)max_leaf) {          // coincident centroids: split by Morton rank
            b.split_bin = -2; b.rank_mid = b.rank_lo + (b.rank_hi - b.rank_lo) / 2u;
            flag = 1;
            SplitChildren c; c.lbox = aabb_empty(); c.rbox = aabb_empty(); c.lcount = 0; c.rcount = 0;   // filled by atomics in k_relabel
            sc[node - lb] = c;
        }
    } else {
        float best_cost = FLT_MAX; int best_bin = -1, best_axis = -1;
        for (int ax = 0; ax < 3; ax++) {
            if (!(b.cscale[ax] > 0.f)) continue;
            const Bin* nb = bins + ((size_t)(node - lb) * 3 + ax) * kBins;
            float right_area[kBins];
            aabb acc = aabb_empty();
            for (int k = kBins - 1; k > 0; k--) {
                if (nb[k].count) aabb_grow(acc, ordbox_to_aabb(nb[k].box));
                right_area[k] = aabb_half_area(acc);
            }
            acc = aabb_empty();
            uint32_t nl = 0;
            for (int k = 0; k < kBins - 1; k++) {
                if (nb[k].count) aabb_grow(acc, ordbox_to_aabb(nb[k].box));
                nl += nb[k].count;
                uint32_t nr = cnt - nl;
                if (nl == 0 || nr == 0) continue;
                float c = aabb_half_area(acc) * (float)nl + right_area[k + 1] * (float)nr;
                if (c < best_cost) { best_cost = c; best_bin = k; best_axis = ax; }
            }
        }
        const Bin* nb = bins + ((size_t)(node - lb) * 3 + (best_axis < 0 ? 0 : best_axis)) * kBins;
        float parent_area = aabb_half_area(nodes[node].box);
        float split_cost = kTravCost + (parent_area > 0.f ? best_cost / parent_area : (float)cnt);
        bool leaf = cnt <= (uint32_t)max_leaf && (float)cnt <= split_cost;
        if (best_bin < 0) {                       // cannot happen with a positive centroid extent; be safe
            leaf = cnt <= (uint32_t)max_leaf;
            if (!leaf) { b.split_bin = -2; b.rank_mid = b.rank_lo + (b.rank_hi - b.rank_lo) / 2u; flag = 1;
                         SplitChildren c; c.lbox = aabb_empty(); c.rbox = aabb_empty(); c.lcount = 0; c.rcount = 0; sc[node - lb] = c; }
        } else if (!leaf) {
            SplitChildren c; c.lbox = aabb_empty(); c.rbox = aabb_empty(); c.lcount = 0; c.rcount = 0;
            for (int k = 0; k < kBins; k++) {
                if (!nb[k].count) continue;
                if (k <= best_bin) { aabb_grow(c.lbox, ordbox_to_aabb(nb[k].box)); c.lcount += nb[k].count; }
                else { aabb_grow(c.rbox, ordbox_to_aabb(nb[k].box)); c.rcount += nb[k].count; }
            }
            sc[node - lb] = c;
            b.split_bin = best_bin; b.axis = best_axis; flag = 1;
        }
    }
    if (!flag) { nodes[node].count = cnt; nodes[node].left = 0; nodes[node].right = 0; }     // leaf; `left` = first slot, set after the final sort
    bn[node] = b;
    flags[node - lb] = flag;
}

__global__ void __launch_bounds__(kThreads) k_make_children(Bvh2Node* nodes, BuildNode* bn, uint32_t lb, uint32_t le, const uint32_t* __restrict__ flags,
                                                            const uint32_t* __restrict__ scan, const SplitChildren* __restrict__ sc) {
    uint32_t node = lb + blockIdx.x * kThreads + threadIdx.x;
    if (node >= le || !flags[node - lb]) return;
    uint32_t l = le + 2u * scan[node - lb];
    nodes[node].left = l; nodes[node].right = l + 1; nodes[node].count = 0;
    const SplitChildren c = sc[node - lb];
    const bool fallback = bn[node].split_bin == -2;
    for (int side = 0; side < 2; side++) {
        Bvh2Node nn; nn.box = side ? c.rbox : c.lbox; nn.left = 0; nn.right = 0; nn.count = 0;
        nodes[l + side] = nn;
        BuildNode b;
        b.cbox = ordbox_empty(); b.rank_lo = 0xFFFFFFFFu; b.rank_hi = 0u; b.axis = -1; for (int a = 0; a < 3; a++) { b.cmin[a] = 0.f; b.cscale[a] = 0.f; } b.split_bin = -1; b.rank_mid = 0;
        b.pcount = fallback ? 0u : (side ? c.rcount : c.lcount);
        b.fbox = ordbox_empty(); b.box_from_atomics = fallback ? 1u : 0u;
        bn[l + side] = b;
    }
}

__global__ void __launch_bounds__(kThreads) k_relabel(const PrimBox* __restrict__ boxes, uint32_t* prim_node, uint32_t n, const Bvh2Node* __restrict__ nodes,
                                                      BuildNode* bn, uint32_t lb, uint32_t le) {
    uint32_t r = blockIdx.x * kThreads + threadIdx.x;
    bool active = r < n;
    uint32_t node = active ? prim_node[r] : 0u;
    active = active && node >= lb && node < le;
    const uint32_t left = active ? nodes[node].left : 0u;
    active = active && left != 0;                 // left == 0: the node became a leaf
    const unsigned live = __ballot_sync(0xFFFFFFFFu, active);
    if (!active) return;
    const PrimBox pb = boxes[r];
    const BuildNode* b = bn + node;
    const int32_t split_bin = b->split_bin;
    uint32_t side;
    if (split_bin == -2) side = r > b->rank_mid ? 1u : 0u;
    else {
        const int ax = b->axis;
        side = bin_of(pb, ax, b->cmin[ax], b->cscale[ax]) > split_bin ? 1u : 0u;
    }
    uint32_t child = left + side;
    prim_node[r] = child;
    BuildNode* cb = bn + child;
    float c[3];
    for (int a = 0; a < 3; a++) c[a] = 0.5f * (pb.lo[a] + pb.hi[a]);
    // one set of atomics per (warp, child) instead of per primitive (see ordbox_group_grow)
    const unsigned group = __match_any_sync(live, child);
    const bool leader = (int)(threadIdx.x & 31) == __ffs(group) - 1;
    ordbox_group_grow(&cb->cbox, group, leader, c, c);
    const uint32_t rlo = __reduce_min_sync(group, r), rhi = __reduce_max_sync(group, r);
    if (leader) { atomicMin(&cb->rank_lo, rlo); atomicMax(&cb->rank_hi, rhi); }
    if (split_bin == -2) {                        // every lane of a group shares the parent, hence split_bin
        ordbox_group_grow(&cb->fbox, group, leader, pb.lo, pb.hi);
        if (leader) atomicAdd(&cb->pcount, (uint32_t)__popc(group));
    }
}

// after the final (leaf id, rank) sort: first slot of every leaf and the triangle order
__global__ void __launch_bounds__(kThreads) k_leaf_first(const uint32_t* __restrict__ leaf_sorted, const uint32_t* __restrict__ rank_sorted,
                                                         const PrimBox* __restrict__ boxes, uint32_t n, Bvh2Node* nodes, uint32_t* tri_order) {
    uint32_t p = blockIdx.x * kThreads + threadIdx.x;
    if (p >= n) return;
    uint32_t leaf = leaf_sorted[p];
    if (p == 0 || leaf_sorted[p - 1] != leaf) nodes[leaf].left = p;
    tri_order[p] = boxes[rank_sorted[p]].tri;
}

// ---------------------------------------------------------------------------------------------
// collapse: one work item = (BVH2 subtree root, index of the BVH8 node it becomes)
// ---------------------------------------------------------------------------------------------
struct WorkItem {
    uint32_t b2node, out_index;
};

__global__ void __launch_bounds__(kThreads) k_collapse_plan(const Bvh2Node* __restrict__ nodes, const WorkItem* __restrict__ items, uint32_t n_items,
                                                            uint32_t* n_inner, uint32_t* n_tris) {
    uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= n_items) return;
    WideNodePlan plan;
    plan_wide_node(nodes, items[i].b2node, plan);
    n_inner[i] = (uint32_t)plan.n_inner;
    n_tris[i] = plan.n_tris;
}

__global__ void __launch_bounds__(kThreads) k_collapse_emit(const Bvh2Node* __restrict__ nodes, const WorkItem* __restrict__ items, uint32_t n_items,
                                                            const uint32_t* __restrict__ inner_off, const uint32_t* __restrict__ tri_off,
                                                            uint32_t node_total, uint32_t tri_total, const TriIn* __restrict__ tris,
                                                            const uint32_t* __restrict__ tri_order, u32x4* out_nodes, f32x4* out_tris,
                                                            WorkItem* next_items, float inv_root_area, float* sah) {
    uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    float cost = 0.f;
    if (i < n_items) {
    const WorkItem it = items[i];
    WideNodePlan plan;
    plan_wide_node(nodes, it.b2node, plan);
    const uint32_t child_base = node_total + inner_off[i];
    const uint32_t tri_base = tri_total + tri_off[i];
    u32x4 enc[5];
    encode_node(nodes[it.b2node].box, plan.slots, child_base, tri_base, enc);
    for (int k = 0; k < 5; k++) out_nodes[(size_t)it.out_index * kNodeQuads + k] = enc[k];
    uint32_t t = tri_base;
    for (int l = 0; l < plan.n_leaf; l++) {
        const Bvh2Node c = nodes[plan.leaf_kids[l]];
        for (uint32_t k = 0; k < c.count; k++, t++) {
            f32x4 rec[3];
            pack_triangle(tris[tri_order[c.left + k]], rec);
            out_tris[(size_t)t * kTriQuads + 0] = rec[0]; out_tris[(size_t)t * kTriQuads + 1] = rec[1]; out_tris[(size_t)t * kTriQuads + 2] = rec[2];
        }
    }
    for (int j = 0; j < plan.n_inner; j++) { WorkItem w; w.b2node = plan.inner_kids[j]; w.out_index = child_base + (uint32_t)j; next_items[inner_off[i] + j] = w; }
    cost = (plan.leaf_area_cost + aabb_half_area(nodes[it.b2node].box) * kTravCost) * inv_root_area;
    }
    // statistics only: warp-reduced float atomic
    for (int o = 16; o > 0; o >>= 1) cost += __shfl_xor_sync(0xFFFFFFFFu, cost, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(sah, cost);
}

__global__ void k_empty_root(u32x4* out_nodes) {
    WideChild slots[8];
    for (int s = 0; s < 8; s++) { slots[s].box = aabb_empty(); slots[s].is_inner = 0; slots[s].tri_count = 0; slots[s].tri_offset = 0; }
    aabb z; z.lo = mk3(0, 0, 0); z.hi = mk3(0, 0, 0);
    u32x4 enc[5];
    encode_node(z, slots, 1, 0, enc);
    for (int k = 0; k < 5; k++) out_nodes[k] = enc[k];
}

__global__ void __launch_bounds__(kThreads) k_iota(uint32_t* v, uint32_t n) {
    uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    if (i < n) v[i] = i;
}

inline unsigned grid_for(size_t n) { return (unsigned)((n + kThreads - 1) / kThreads); }

struct DevBuf {            // frees on scope exit
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { if (p) { cudaFree(p); p = nullptr; } return cudaMalloc(&p, bytes ? bytes : 16); }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace

#define BG_CUDA(call)                                                                   \
    do {                                                                                \
        cudaError_t e_ = (call);                                                        \
        if (e_ != cudaSuccess) { err = std::string(#call) + ": " + cudaGetErrorString(e_); return e_; } \
    } while (0)

cudaError_t build_bvh8_gpu(cudaStream_t st, const TriIn* tris, size_t n_in, int max_leaf, int max_bvh8_depth, GpuBvhResult& out,
                           std::string& err, bool tris_on_device) {
    out = GpuBvhResult();
    if (max_leaf < 1 || max_leaf > 3) max_leaf = 3;
    const uint32_t n = (uint32_t)n_in;
    if (n == 0) {                                             // empty scene: a root with no children
        BG_CUDA(cudaMalloc(&out.d_nodes, sizeof(u32x4) * kNodeQuads));
        BG_CUDA(cudaMalloc(&out.d_tris, sizeof(f32x4) * kTriQuads));
        k_empty_root<<<1, 1, 0, st>>>(out.d_nodes);
        BG_CUDA(cudaGetLastError());
        BG_CUDA(cudaStreamSynchronize(st));
        out.n_nodes = 1; out.max_depth = 1;
        out.bounds.lo = mk3(0, 0, 0); out.bounds.hi = mk3(0, 0, 0);
        return cudaSuccess;
    }
    cudaEvent_t ev[5];
    for (auto& e : ev) BG_CUDA(cudaEventCreate(&e));
    struct EvGuard { cudaEvent_t* e; ~EvGuard() { for (int i = 0; i < 5; i++) cudaEventDestroy(e[i]); } } ev_guard{ev};
    BG_CUDA(cudaEventRecord(ev[0], st));
    const size_t node_cap = 2 * (size_t)n;                    // a binary tree over n primitives has < 2n nodes
    DevBuf d_tris_in, d_boxes0, d_boxes, d_scene, d_keys, d_keys2, d_vals, d_vals2, d_nodes, d_bn, d_prim_node, d_flags, d_scan, d_sc, d_bins, d_tmp;
    const TriIn* d_in = tris;                                 // device-resident input (geometry_gpu.cu) is used in place
    if (!tris_on_device) {
        BG_CUDA(d_tris_in.alloc(sizeof(TriIn) * n));
        BG_CUDA(cudaMemcpyAsync(d_tris_in.p, tris, sizeof(TriIn) * n, cudaMemcpyHostToDevice, st));
        d_in = d_tris_in.as<TriIn>();
    }
    BG_CUDA(d_boxes0.alloc(sizeof(PrimBox) * n));
    BG_CUDA(d_boxes.alloc(sizeof(PrimBox) * n));
    BG_CUDA(d_scene.alloc(sizeof(OrdBox) * 2));
    BG_CUDA(cudaMemsetAsync(d_scene.p, 0, sizeof(OrdBox) * 2, st));      // all-zero = empty box
    k_prim_setup<<<grid_for(n), kThreads, 0, st>>>(d_in, n, d_boxes0.as<PrimBox>(), d_scene.as<OrdBox>());
    // ---- Morton order ------------------------------------------------------------------------
    BG_CUDA(d_keys.alloc(4 * (size_t)n)); BG_CUDA(d_keys2.alloc(4 * (size_t)n)); BG_CUDA(d_vals.alloc(4 * (size_t)n)); BG_CUDA(d_vals2.alloc(4 * (size_t)n));
    k_morton<<<grid_for(n), kThreads, 0, st>>>(d_boxes0.as<PrimBox>(), n, d_scene.as<OrdBox>(), d_keys.as<uint32_t>(), d_vals.as<uint32_t>());
    size_t tmp_bytes = 0, tmp_scan = 0;
    BG_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_keys.as<uint32_t>(), d_keys2.as<uint32_t>(), d_vals.as<uint32_t>(), d_vals2.as<uint32_t>(), (int)n, 0, 32, st));
    BG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan, d_keys.as<uint32_t>(), d_keys2.as<uint32_t>(), (int)n, st));
    tmp_bytes = std::max(tmp_bytes, tmp_scan);
    BG_CUDA(d_tmp.alloc(tmp_bytes));
    BG_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, tmp_bytes, d_keys.as<uint32_t>(), d_keys2.as<uint32_t>(), d_vals.as<uint32_t>(), d_vals2.as<uint32_t>(), (int)n, 0, 30, st));
    k_gather_boxes<<<grid_for(n), kThreads, 0, st>>>(d_boxes0.as<PrimBox>(), d_vals2.as<uint32_t>(), n, d_boxes.as<PrimBox>());
    BG_CUDA(cudaGetLastError());
    BG_CUDA(cudaEventRecord(ev[1], st));
    // ---- breadth-first binned SAH ---------------------------------------------------------------
    BG_CUDA(d_nodes.alloc(sizeof(Bvh2Node) * node_cap));
    BG_CUDA(d_bn.alloc(sizeof(BuildNode) * node_cap));
    BG_CUDA(d_prim_node.alloc(4 * (size_t)n));
    BG_CUDA(cudaMemsetAsync(d_prim_node.p, 0, 4 * (size_t)n, st));
    k_init_root<<<1, 1, 0, st>>>(d_nodes.as<Bvh2Node>(), d_bn.as<BuildNode>(), d_scene.as<OrdBox>(), n);
    size_t level_cap = 0;
    uint32_t lb = 0, le = 1;
    int levels = 0;
    while (le > lb) {
        if (++levels > 128) { err = "binned-SAH build exceeded 128 levels"; return cudaErrorInvalidValue; }
        const uint32_t width = le - lb;
        if (width > level_cap) {
            level_cap = std::max<size_t>((size_t)width * 2, 1024);
            BG_CUDA(cudaStreamSynchronize(st));
            BG_CUDA(d_flags.alloc(4 * level_cap)); BG_CUDA(d_scan.alloc(4 * level_cap));
            BG_CUDA(d_sc.alloc(sizeof(SplitChildren) * level_cap)); BG_CUDA(d_bins.alloc(sizeof(Bin) * 3 * kBins * level_cap));
        }
        k_prepare<<<grid_for(width), kThreads, 0, st>>>(d_nodes.as<Bvh2Node>(), d_bn.as<BuildNode>(), lb, le);
        // all-zero = empty boxes, count 0
        BG_CUDA(cudaMemsetAsync(d_bins.p, 0, sizeof(Bin) * 3 * kBins * (size_t)width, st));
        k_bin<<<grid_for(n), kThreads, 0, st>>>(d_boxes.as<PrimBox>(), d_prim_node.as<uint32_t>(), n, d_bn.as<BuildNode>(), lb, le, d_bins.as<Bin>());
        k_split<<<grid_for(width), kThreads, 0, st>>>(d_nodes.as<Bvh2Node>(), d_bn.as<BuildNode>(), lb, le, d_bins.as<Bin>(), max_leaf,
                                                       d_flags.as<uint32_t>(), d_sc.as<SplitChildren>());
        BG_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tmp_bytes, d_flags.as<uint32_t>(), d_scan.as<uint32_t>(), (int)width, st));
        k_make_children<<<grid_for(width), kThreads, 0, st>>>(d_nodes.as<Bvh2Node>(), d_bn.as<BuildNode>(), lb, le, d_flags.as<uint32_t>(),
                                                               d_scan.as<uint32_t>(), d_sc.as<SplitChildren>());
        k_relabel<<<grid_for(n), kThreads, 0, st>>>(d_boxes.as<PrimBox>(), d_prim_node.as<uint32_t>(), n, d_nodes.as<Bvh2Node>(), d_bn.as<BuildNode>(), lb, le);
        BG_CUDA(cudaGetLastError());
        uint32_t last[2];
        BG_CUDA(cudaMemcpyAsync(&last[0], d_scan.as<uint32_t>() + (width - 1), 4, cudaMemcpyDeviceToHost, st));
        BG_CUDA(cudaMemcpyAsync(&last[1], d_flags.as<uint32_t>() + (width - 1), 4, cudaMemcpyDeviceToHost, st));
        BG_CUDA(cudaStreamSynchronize(st));
        const uint32_t nsplit = last[0] + last[1];
        lb = le; le = le + 2 * nsplit;
        if (le > node_cap) { err = "binned-SAH build: node capacity exceeded"; return cudaErrorInvalidValue; }
    }
    const uint32_t n_b2 = le;
    out.n_bvh2_nodes = n_b2; out.bvh2_levels = levels;
    BG_CUDA(cudaEventRecord(ev[2], st));
    // ---- leaf order: stable sort of (leaf id, rank) -----------------------------------------------------
    DevBuf d_rank, d_leaf_sorted, d_rank_sorted, d_tri_order;
    BG_CUDA(d_rank.alloc(4 * (size_t)n)); BG_CUDA(d_leaf_sorted.alloc(4 * (size_t)n)); BG_CUDA(d_rank_sorted.alloc(4 * (size_t)n));
    BG_CUDA(d_tri_order.alloc(4 * (size_t)n));
    k_iota<<<grid_for(n), kThreads, 0, st>>>(d_rank.as<uint32_t>(), n);
    {
        int bits = 1;
        while ((1ull << bits) < (unsigned long long)n_b2) bits++;
        size_t need = 0;
        BG_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, d_prim_node.as<uint32_t>(), d_leaf_sorted.as<uint32_t>(), d_rank.as<uint32_t>(),
                                                d_rank_sorted.as<uint32_t>(), (int)n, 0, bits, st));
        if (need > tmp_bytes) { BG_CUDA(cudaStreamSynchronize(st)); tmp_bytes = need; BG_CUDA(d_tmp.alloc(tmp_bytes)); }
        BG_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, tmp_bytes, d_prim_node.as<uint32_t>(), d_leaf_sorted.as<uint32_t>(), d_rank.as<uint32_t>(),
                                                d_rank_sorted.as<uint32_t>(), (int)n, 0, bits, st));
    }
    k_leaf_first<<<grid_for(n), kThreads, 0, st>>>(d_leaf_sorted.as<uint32_t>(), d_rank_sorted.as<uint32_t>(), d_boxes.as<PrimBox>(), n,
                                                    d_nodes.as<Bvh2Node>(), d_tri_order.as<uint32_t>());
    BG_CUDA(cudaGetLastError());
    BG_CUDA(cudaEventRecord(ev[3], st));
    // ---- collapse to BVH8, level by level -----------------------------------------------------------------
    const size_t n_b2_inner = (n_b2 - 1) / 2;                 // a full binary tree: inner = (nodes - 1) / 2
    const size_t out_cap = n_b2_inner + 1;                    // every BVH8 node replaces >= 1 BVH2 inner node (or is the lone root)
    DevBuf d_out_nodes, d_items[2], d_ninner, d_ntris, d_ioff, d_toff, d_sah;
    BG_CUDA(d_out_nodes.alloc(sizeof(u32x4) * kNodeQuads * out_cap));
    BG_CUDA(cudaMalloc(&out.d_tris, sizeof(f32x4) * kTriQuads * (size_t)n));
    BG_CUDA(d_items[0].alloc(sizeof(WorkItem) * out_cap)); BG_CUDA(d_items[1].alloc(sizeof(WorkItem) * out_cap));
    BG_CUDA(d_ninner.alloc(4 * out_cap)); BG_CUDA(d_ntris.alloc(4 * out_cap)); BG_CUDA(d_ioff.alloc(4 * out_cap)); BG_CUDA(d_toff.alloc(4 * out_cap));
    BG_CUDA(d_sah.alloc(4)); BG_CUDA(cudaMemsetAsync(d_sah.p, 0, 4, st));
    {
        size_t need = 0;
        BG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, d_ninner.as<uint32_t>(), d_ioff.as<uint32_t>(), (int)out_cap, st));
        if (need > tmp_bytes) { BG_CUDA(cudaStreamSynchronize(st)); tmp_bytes = need; BG_CUDA(d_tmp.alloc(tmp_bytes)); }
    }
    Bvh2Node root;
    BG_CUDA(cudaMemcpyAsync(&root, d_nodes.p, sizeof root, cudaMemcpyDeviceToHost, st));
    WorkItem first{0u, 0u};
    BG_CUDA(cudaMemcpyAsync(d_items[0].p, &first, sizeof first, cudaMemcpyHostToDevice, st));
    BG_CUDA(cudaStreamSynchronize(st));
    out.bounds = root.box;
    float root_area = aabb_half_area(root.box);
    if (!(root_area > 0.f)) root_area = 1.f;
    uint32_t n_items = 1, node_total = 1, tri_total = 0;
    int cur = 0, depth = 0;
    while (n_items > 0) {
        depth++;
        if (depth > max_bvh8_depth) {
            cudaFree(out.d_tris); out.d_tris = nullptr;
            err = "BVH8 deeper than the traversal stack (" + std::to_string(max_bvh8_depth) + " levels)";
            return cudaErrorInvalidValue;
        }
        k_collapse_plan<<<grid_for(n_items), kThreads, 0, st>>>(d_nodes.as<Bvh2Node>(), d_items[cur].as<WorkItem>(), n_items, d_ninner.as<uint32_t>(),
                                                                 d_ntris.as<uint32_t>());
        BG_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tmp_bytes, d_ninner.as<uint32_t>(), d_ioff.as<uint32_t>(), (int)n_items, st));
        BG_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tmp_bytes, d_ntris.as<uint32_t>(), d_toff.as<uint32_t>(), (int)n_items, st));
        uint32_t tail[4];
        BG_CUDA(cudaMemcpyAsync(&tail[0], d_ioff.as<uint32_t>() + (n_items - 1), 4, cudaMemcpyDeviceToHost, st));
        BG_CUDA(cudaMemcpyAsync(&tail[1], d_ninner.as<uint32_t>() + (n_items - 1), 4, cudaMemcpyDeviceToHost, st));
        BG_CUDA(cudaMemcpyAsync(&tail[2], d_toff.as<uint32_t>() + (n_items - 1), 4, cudaMemcpyDeviceToHost, st));
        BG_CUDA(cudaMemcpyAsync(&tail[3], d_ntris.as<uint32_t>() + (n_items - 1), 4, cudaMemcpyDeviceToHost, st));
        BG_CUDA(cudaStreamSynchronize(st));
        const uint32_t next_items = tail[0] + tail[1], level_tris = tail[2] + tail[3];
        if ((size_t)node_total + next_items > out_cap || (size_t)tri_total + level_tris > n) {
            cudaFree(out.d_tris); out.d_tris = nullptr;
            err = "BVH8 collapse: capacity exceeded"; return cudaErrorInvalidValue;
        }
        k_collapse_emit<<<grid_for(n_items), kThreads, 0, st>>>(d_nodes.as<Bvh2Node>(), d_items[cur].as<WorkItem>(), n_items, d_ioff.as<uint32_t>(),
                                                                 d_toff.as<uint32_t>(), node_total, tri_total, d_in, d_tri_order.as<uint32_t>(),
                                                                 d_out_nodes.as<u32x4>(), out.d_tris, d_items[cur ^ 1].as<WorkItem>(), 1.0f / root_area,
                                                                 d_sah.as<float>());
        BG_CUDA(cudaGetLastError());
        node_total += next_items; tri_total += level_tris;
        n_items = next_items; cur ^= 1;
    }
    // exact-size node array
    BG_CUDA(cudaMalloc(&out.d_nodes, sizeof(u32x4) * kNodeQuads * (size_t)node_total));
    BG_CUDA(cudaMemcpyAsync(out.d_nodes, d_out_nodes.p, sizeof(u32x4) * kNodeQuads * (size_t)node_total, cudaMemcpyDeviceToDevice, st));
    BG_CUDA(cudaMemcpyAsync(&out.sah_cost, d_sah.p, 4, cudaMemcpyDeviceToHost, st));
    BG_CUDA(cudaEventRecord(ev[4], st));
    BG_CUDA(cudaStreamSynchronize(st));
    for (int i = 0; i < 4; i++) cudaEventElapsedTime(&out.stage_ms[i], ev[i], ev[i + 1]);
    out.n_nodes = node_total; out.n_tris = tri_total; out.max_depth = depth;
    if (tri_total != n) { err = "BVH8 collapse lost triangles"; return cudaErrorInvalidValue; }
    return cudaSuccess;
}

}  // namespace ptx
