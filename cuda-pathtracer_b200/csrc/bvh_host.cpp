// bvh_host.cpp -- see bvh_host.h.  Host binned-SAH BVH2, greedy surface-area collapse to BVH8,
// octant-aware slot assignment and 80-byte node compression (layout.h).
#include "bvh_host.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace ptx {

namespace {

struct BuildItem {
    uint32_t node, begin, end;
};

inline aabb tri_box(const TriIn& t) {
    aabb b = aabb_empty();
    aabb_grow(b, ld3(t.v0)); aabb_grow(b, ld3(t.v1)); aabb_grow(b, ld3(t.v2));
    return b;
}
inline vec3 box_center(const aabb& b) { return mk3(0.5f * (b.lo.x + b.hi.x), 0.5f * (b.lo.y + b.hi.y), 0.5f * (b.lo.z + b.hi.z)); }
inline aabb padded(const aabb& b) {     // bvh_common.h box_pad; the centroid is taken from the unpadded box (as k_prim_setup does)
    float px = box_pad(b.lo.x, b.hi.x), py = box_pad(b.lo.y, b.hi.y), pz = box_pad(b.lo.z, b.hi.z);
    aabb r;
    r.lo = mk3(b.lo.x - px, b.lo.y - py, b.lo.z - pz);
    r.hi = mk3(b.hi.x + px, b.hi.y + py, b.hi.z + pz);
    return r;
}


}  // namespace

void build_bvh2_sah(const TriIn* tris, size_t n, int bins, int max_leaf, Bvh2& out) {
    if (bins <= 1 || bins > 64) bins = 16;
    if (max_leaf < 1 || max_leaf > 3) max_leaf = 3;
    out.nodes.clear(); out.order.resize(n);
    std::vector<aabb> boxes(n);
    std::vector<vec3> cents(n);
    for (size_t i = 0; i < n; i++) { out.order[i] = (uint32_t)i; aabb tb = tri_box(tris[i]); cents[i] = box_center(tb); boxes[i] = padded(tb); }
    out.nodes.reserve(n ? 2 * n : 1);
    out.nodes.push_back(Bvh2Node{aabb_empty(), 0, 0, 0});
    if (n == 0) return;

    std::vector<BuildItem> todo;
    todo.push_back({0u, 0u, (uint32_t)n});
    std::vector<aabb> bin_box((size_t)bins);
    std::vector<uint32_t> bin_cnt((size_t)bins);
    std::vector<float> right_area((size_t)bins);

    while (!todo.empty()) {
        BuildItem it = todo.back(); todo.pop_back();
        uint32_t cnt = it.end - it.begin;
        aabb nb = aabb_empty(), cb = aabb_empty();
        for (uint32_t i = it.begin; i < it.end; i++) { uint32_t t = out.order[i]; aabb_grow(nb, boxes[t]); aabb_grow(cb, cents[t]); }
        out.nodes[it.node].box = nb;
        auto make_leaf = [&]() { out.nodes[it.node].left = it.begin; out.nodes[it.node].right = 0; out.nodes[it.node].count = cnt; };
        if (cnt == 1) { make_leaf(); continue; }

        // binned SAH over the three axes
        float best_cost = FLT_MAX; int best_axis = -1, best_bin = -1;
        float parent_area = aabb_half_area(nb);
        for (int axis = 0; axis < 3; axis++) {
            float lo = comp(cb.lo, axis), hi = comp(cb.hi, axis);
            if (!(hi > lo)) continue;
            float scale = (float)bins / (hi - lo);
            for (int b = 0; b < bins; b++) { bin_box[b] = aabb_empty(); bin_cnt[b] = 0; }
            for (uint32_t i = it.begin; i < it.end; i++) {
                uint32_t t = out.order[i];
                int b = (int)((comp(cents[t], axis) - lo) * scale);
                b = b < 0 ? 0 : (b >= bins ? bins - 1 : b);
                aabb_grow(bin_box[b], boxes[t]); bin_cnt[b]++;
            }
            aabb acc = aabb_empty();
            for (int b = bins - 1; b > 0; b--) { aabb_grow(acc, bin_box[b]); right_area[b] = aabb_half_area(acc); }
            acc = aabb_empty(); uint32_t nl = 0;
            for (int b = 0; b < bins - 1; b++) {
                aabb_grow(acc, bin_box[b]); nl += bin_cnt[b];
                uint32_t nr = cnt - nl;
                if (nl == 0 || nr == 0) continue;
                float c = aabb_half_area(acc) * (float)nl + right_area[b + 1] * (float)nr;
                if (c < best_cost) { best_cost = c; best_axis = axis; best_bin = b; }
            }
        }
        uint32_t mid;
        if (best_axis < 0) {
            if (cnt <= (uint32_t)max_leaf) { make_leaf(); continue; }
            mid = it.begin + cnt / 2;                     // coincident centroids: split by index
        } else {
            float split_cost = kTravCost + (parent_area > 0.f ? best_cost / parent_area : (float)cnt);
            if (cnt <= (uint32_t)max_leaf && (float)cnt <= split_cost) { make_leaf(); continue; }
            float lo = comp(cb.lo, best_axis), hi = comp(cb.hi, best_axis);
            float scale = (float)bins / (hi - lo);
            auto first = out.order.begin() + it.begin, last = out.order.begin() + it.end;
            auto m = std::partition(first, last, [&](uint32_t t) {
                int b = (int)((comp(cents[t], best_axis) - lo) * scale);
                b = b < 0 ? 0 : (b >= bins ? bins - 1 : b);
                return b <= best_bin;
            });
            mid = (uint32_t)(m - out.order.begin());
            if (mid == it.begin || mid == it.end) mid = it.begin + cnt / 2;
        }
        uint32_t l = (uint32_t)out.nodes.size();
        out.nodes.push_back(Bvh2Node{aabb_empty(), 0, 0, 0});
        out.nodes.push_back(Bvh2Node{aabb_empty(), 0, 0, 0});
        out.nodes[it.node].left = l; out.nodes[it.node].right = l + 1; out.nodes[it.node].count = 0;
        todo.push_back({l + 1, mid, it.end});
        todo.push_back({l, it.begin, mid});
    }
}

// Greedy surface-area collapse (open the largest inner child until 8 children), octant-aware slot
// assignment (slot s = corner (x:bit2, y:bit1, z:bit0) of the node the child sits in, so that visiting
// slots by descending (s ^ octinv) is approximately front to back), breadth-first emission so that a
// node's inner children and its triangles are contiguous.
void collapse_bvh8(const Bvh2& b2, const TriIn* tris, size_t n, Bvh8Result& out) {
    out.nodes.clear(); out.tris.clear(); out.tri_src.clear();
    out.sah_cost = 0.f; out.max_depth = 0;
    out.nodes.resize(kNodeQuads);
    out.tris.reserve(n * kTriQuads); out.tri_src.reserve(n);
    WideChild empty_slots[8];
    for (auto& w : empty_slots) { w.box = aabb_empty(); w.is_inner = 0; w.tri_count = 0; w.tri_offset = 0; }
    if (n == 0 || b2.nodes.empty()) {
        aabb z; z.lo = mk3(0, 0, 0); z.hi = mk3(0, 0, 0);
        out.bounds = z;
        encode_node(z, empty_slots, 1, 0, out.nodes.data());
        return;
    }
    out.bounds = b2.nodes[0].box;
    float root_area = aabb_half_area(out.bounds);
    if (!(root_area > 0.f)) root_area = 1.f;

    struct Work { uint32_t b2node, out_index; int depth; };
    std::vector<Work> queue;
    queue.push_back({0u, 0u, 1});
    size_t head = 0;
    while (head < queue.size()) {
        Work w = queue[head++];
        out.max_depth = std::max(out.max_depth, w.depth);
        const Bvh2Node& nd = b2.nodes[w.b2node];
        WideNodePlan plan;
        plan_wide_node(b2.nodes.data(), w.b2node, plan);
        uint32_t child_base = (uint32_t)(out.nodes.size() / kNodeQuads);
        uint32_t tri_base = (uint32_t)out.tri_src.size();
        uint32_t n_inner = (uint32_t)plan.n_inner;
        for (int l = 0; l < plan.n_leaf; l++) {
            const Bvh2Node& c = b2.nodes[plan.leaf_kids[l]];
            for (uint32_t k = 0; k < c.count; k++) {
                uint32_t src = b2.order[c.left + k];
                f32x4 rec[3]; pack_triangle(tris[src], rec);
                out.tris.push_back(rec[0]); out.tris.push_back(rec[1]); out.tris.push_back(rec[2]);
                out.tri_src.push_back(src);
            }
        }
        out.sah_cost += plan.leaf_area_cost / root_area;
        for (int i = 0; i < plan.n_inner; i++) queue.push_back({plan.inner_kids[i], child_base + (uint32_t)i, w.depth + 1});
        const WideChild* slots = plan.slots;
        out.nodes.resize(out.nodes.size() + (size_t)n_inner * kNodeQuads);
        encode_node(nd.box, slots, child_base, tri_base, out.nodes.data() + (size_t)w.out_index * kNodeQuads);
        out.sah_cost += aabb_half_area(nd.box) / root_area * kTravCost;
    }
}

}  // namespace ptx
