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

constexpr float kTravCost = 1.0f;   // cost of visiting one wide-node child box relative to one triangle test

}  // namespace

void build_bvh2_sah(const TriIn* tris, size_t n, int bins, int max_leaf, Bvh2& out) {
    if (bins <= 1 || bins > 64) bins = 16;
    if (max_leaf < 1 || max_leaf > 3) max_leaf = 3;
    out.nodes.clear(); out.order.resize(n);
    std::vector<aabb> boxes(n);
    std::vector<vec3> cents(n);
    for (size_t i = 0; i < n; i++) { out.order[i] = (uint32_t)i; boxes[i] = tri_box(tris[i]); cents[i] = box_center(boxes[i]); }
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

void pack_triangle(const TriIn& t, f32x4 out[3]) {
    vec3 v0 = ld3(t.v0), e1 = ld3(t.v1) - v0, e2 = ld3(t.v2) - v0;
    out[0] = f32x4{v0.x, v0.y, v0.z, u2f(t.prim_id)};
    out[1] = f32x4{e1.x, e1.y, e1.z, 0.f};
    out[2] = f32x4{e2.x, e2.y, e2.z, 0.f};
}

// Quantisation frame + child boxes.  Exponent e (biased byte) is the smallest with
// ceil((hi - lo) / 2^(e-127)) <= 255; lo is floored and hi ceiled so decoded boxes contain the originals.
void encode_node(const aabb& nb, const WideChild* slots, uint32_t child_base, uint32_t tri_base, u32x4 out[5]) {
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
            // guard against rounding in the division above
            while (eb < 254 && std::ceil(ext[a] / u2f(eb << 23)) > 255.0f) eb++;
        }
        e[a] = eb;
        inv_scale[a] = eb ? 1.0f / u2f(eb << 23) : 0.f;
    }
    uint8_t meta[8], q[6][8];
    uint32_t imask = 0;
    for (int s = 0; s < 8; s++) {
        const WideChild& c = slots[s];
        bool empty = c.box.lo.x > c.box.hi.x;
        if (empty) { meta[s] = 0; for (int k = 0; k < 6; k++) q[k][s] = 0; continue; }
        if (c.is_inner) { meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s)); imask |= 1u << s; }
        else { uint32_t unary = c.tri_count == 1 ? 1u : (c.tri_count == 2 ? 3u : 7u); meta[s] = (uint8_t)((unary << 5) | (c.tri_offset & 31u)); }
        float lo[3] = {c.box.lo.x, c.box.lo.y, c.box.lo.z}, hi[3] = {c.box.hi.x, c.box.hi.y, c.box.hi.z};
        for (int a = 0; a < 3; a++) {
            float scale = e[a] ? u2f(e[a] << 23) : 0.f;
            float ql = std::floor((lo[a] - p[a]) * inv_scale[a]);
            float qh = std::ceil((hi[a] - p[a]) * inv_scale[a]);
            ql = ql < 0.f ? 0.f : (ql > 255.f ? 255.f : ql);
            qh = qh < 0.f ? 0.f : (qh > 255.f ? 255.f : qh);
            // make the decoded box conservative under float rounding of p + q*scale
            while (ql > 0.f && p[a] + ql * scale > lo[a]) ql -= 1.f;
            while (qh < 255.f && p[a] + qh * scale < hi[a]) qh += 1.f;
            q[a][s] = (uint8_t)ql; q[3 + a][s] = (uint8_t)qh;
        }
    }
    auto pack4 = [](const uint8_t* b) { return (uint32_t)b[0] | ((uint32_t)b[1] << 8) | ((uint32_t)b[2] << 16) | ((uint32_t)b[3] << 24); };
    out[0] = u32x4{f2u(p[0]), f2u(p[1]), f2u(p[2]), e[0] | (e[1] << 8) | (e[2] << 16) | (imask << 24)};
    out[1] = u32x4{child_base, tri_base, pack4(meta), pack4(meta + 4)};
    out[2] = u32x4{pack4(q[0]), pack4(q[0] + 4), pack4(q[1]), pack4(q[1] + 4)};
    out[3] = u32x4{pack4(q[2]), pack4(q[2] + 4), pack4(q[3]), pack4(q[3] + 4)};
    out[4] = u32x4{pack4(q[4]), pack4(q[4] + 4), pack4(q[5]), pack4(q[5] + 4)};
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
        uint32_t kids[8]; int nk = 0;
        if (nd.count > 0) kids[nk++] = w.b2node;                       // degenerate: root itself is a leaf
        else { kids[nk++] = nd.left; kids[nk++] = nd.right; }
        while (nk < 8) {
            int pick = -1; float best = -1.f;
            for (int i = 0; i < nk; i++) {
                const Bvh2Node& c = b2.nodes[kids[i]];
                if (c.count > 0) continue;
                float a = aabb_half_area(c.box);
                if (a > best) { best = a; pick = i; }
            }
            if (pick < 0) break;
            uint32_t open = kids[pick];
            kids[pick] = b2.nodes[open].left;
            kids[nk++] = b2.nodes[open].right;
        }
        // slot assignment: greedy maximum of dot(child centre - node centre, corner direction)
        vec3 nc = mk3(0.5f * (nd.box.lo.x + nd.box.hi.x), 0.5f * (nd.box.lo.y + nd.box.hi.y), 0.5f * (nd.box.lo.z + nd.box.hi.z));
        float cost[8][8];
        for (int i = 0; i < nk; i++) {
            const aabb& cb = b2.nodes[kids[i]].box;
            vec3 dlt = mk3(0.5f * (cb.lo.x + cb.hi.x), 0.5f * (cb.lo.y + cb.hi.y), 0.5f * (cb.lo.z + cb.hi.z)) - nc;
            for (int s = 0; s < 8; s++)
                cost[i][s] = ((s & 4) ? dlt.x : -dlt.x) + ((s & 2) ? dlt.y : -dlt.y) + ((s & 1) ? dlt.z : -dlt.z);
        }
        int slot_of[8]; int child_in_slot[8];
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
        WideChild slots[8];
        uint32_t child_base = (uint32_t)(out.nodes.size() / kNodeQuads);
        uint32_t tri_base = (uint32_t)out.tri_src.size();
        uint32_t n_inner = 0, tri_off = 0;
        for (int s = 0; s < 8; s++) {
            slots[s] = empty_slots[0];
            int ci = child_in_slot[s];
            if (ci < 0) continue;
            const Bvh2Node& c = b2.nodes[kids[ci]];
            slots[s].box = c.box;
            if (c.count > 0) {
                slots[s].is_inner = 0; slots[s].tri_count = c.count; slots[s].tri_offset = tri_off;
                for (uint32_t k = 0; k < c.count; k++) {
                    uint32_t src = b2.order[c.left + k];
                    f32x4 rec[3]; pack_triangle(tris[src], rec);
                    out.tris.push_back(rec[0]); out.tris.push_back(rec[1]); out.tris.push_back(rec[2]);
                    out.tri_src.push_back(src);
                }
                tri_off += c.count;
                out.sah_cost += aabb_half_area(c.box) / root_area * (float)c.count;
            } else {
                slots[s].is_inner = 1;
                queue.push_back({kids[ci], child_base + n_inner, w.depth + 1});
                n_inner++;
            }
        }
        out.nodes.resize(out.nodes.size() + (size_t)n_inner * kNodeQuads);
        encode_node(nd.box, slots, child_base, tri_base, out.nodes.data() + (size_t)w.out_index * kNodeQuads);
        out.sah_cost += aabb_half_area(nd.box) / root_area * kTravCost;
    }
}

}  // namespace ptx
