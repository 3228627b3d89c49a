// traverse_inst.h -- two-level traversal: a TLAS over object instances, one BLAS per object (both compressed BVH8, traverse.h).
//
// The reference's loadPBRT flattens every ObjectInstance into world-space triangles (utils/pbrt_loader.h:197 makeSingleLevel,
// :282-344 per-instance expansion): a scene of 305 instances of one 327 680-triangle mesh becomes 100 M triangles -- 4.8 GB of
// records, 1.5 GB of nodes, 30 s of loading.  With ptx_load_opts.keep_instances the object is stored ONCE in object space and the
// ray is transformed at the instance leaf instead.  Primitive NUMBERING stays the flattened order
// (instance.prim_base + index within the object), so ptx_trace_primary_ids reports what the flattened scene reports.
// Hit distances are computed in object space with the UN-normalised transformed direction, so t is the world-space parameter;
// they agree with the flattened scene's to rounding, not bit for bit (documented: DESIGN.md, instancing).
//
// Layout: nodes / tris hold the TLAS first (root = node 0), then every BLAS with its child / triangle bases already offset.
// A TLAS leaf "triangle" is a placeholder record whose w word is the instance index.  Stack discipline: entering an instance
// pushes a marker entry { list base, remaining instance bits } whose top byte is 0; node-group entries always have a hit bit in
// their top byte, so popping a marker means "this BLAS is finished: back to world space".
#pragma once
#include "traverse.h"

namespace ptx {

struct InstanceRec {          // 128 bytes
    float m[12];              // object -> world: columns vx, vy, vz of the linear part, then the translation (pbrtParser affine3f)
    float inv[12];            // world -> object, same layout
    uint32_t root;            // node index of the object's BLAS root
    uint32_t prim_base;       // flattened primitive id of the object's first primitive in THIS instance
    uint32_t shade_base;      // index of the object's first record in prim_shade
    uint32_t flags;           // reserved
    uint32_t pad[4];
};
static_assert(sizeof(InstanceRec) == 128, "InstanceRec is read as eight 16-byte quads");

PTX_HD vec3 xfm_vec(const float* m, vec3 v) {          // linear part only
    return mk3(m[0] * v.x + m[3] * v.y + m[6] * v.z, m[1] * v.x + m[4] * v.y + m[7] * v.z, m[2] * v.x + m[5] * v.y + m[8] * v.z);
}
PTX_HD vec3 xfm_pt(const float* m, vec3 p) {
    return mk3(m[0] * p.x + m[3] * p.y + m[6] * p.z + m[9], m[1] * p.x + m[4] * p.y + m[7] * p.z + m[10], m[2] * p.x + m[5] * p.y + m[8] * p.z + m[11]);
}

struct InstTraversal {
    Traversal tr;             // cursor; closest / best / bestw as in the single-level traversal (bestw = primitive index WITHIN the object)
    vec3 o, d;                // the ray in the space being traversed (world, or the current instance's object space)
    uint32_t inst;            // instance being traversed, kMiss at the top level
    uint32_t best_inst;       // instance of the closest accepted hit

    PTX_HD void begin(vec3 ow, vec3 dw, float t_max) {
        tr.begin(dw, t_max);
        o = ow; d = dw; inst = kMiss; best_inst = kMiss;
    }
    PTX_HD void set_dir(vec3 dir) {                     // 1 / d of Traversal::begin without touching the cursor
        const float ooeps = 8.271806e-25f;
        tr.idx = rcp_fast(fabsf(dir.x) > ooeps ? dir.x : (dir.x < 0.0f ? -ooeps : ooeps));
        tr.idy = rcp_fast(fabsf(dir.y) > ooeps ? dir.y : (dir.y < 0.0f ? -ooeps : ooeps));
        tr.idz = rcp_fast(fabsf(dir.z) > ooeps ? dir.z : (dir.z < 0.0f ? -ooeps : ooeps));
    }

    // One node visit (TLAS or BLAS) + the triangles of the leaves it hit (BLAS) or the instance list it produced (TLAS).
    // Returns true when the traversal is complete.  (ow, dw) = the world-space ray.
    template <int STACK>
    PTX_HD bool step(const SceneView& s, TravStack<STACK>& stk, vec3 ow, vec3 dw, float t_min) {
        uint32_t tg_x, tg_y;
        tr.template visit<false>(s, stk, o, d, t_min, tg_x, tg_y, nullptr);
        // visit() leaves the node's imask in the low byte of ng_y; without a hit inner child the group is exhausted, and below a
        // value <= 0x00FFFFFF with low bits set means "instance list": clear it
        if (tr.ng_y <= 0x00FFFFFFu) tr.ng_y = 0u;
        if (inst == kMiss) {
            if (tg_y != 0) {                            // instances hit by this TLAS node: they become the current group
                if (tr.ng_y > 0x00FFFFFFu) { stk.x[tr.sp] = tr.ng_x; stk.y[tr.sp] = tr.ng_y; tr.sp++; }
                tr.ng_x = tg_x; tr.ng_y = tg_y;
            }
        } else {
            // two records per round, both tested against the distance the round started with, then committed in order: the same
            // accepted hits as one at a time (Traversal::step, PAIR).  ncu r2x on C5: the one-record loop ran 3.2 rounds per
            // iteration at lane efficiency 0.09, 21 % of the kernel's warp instructions.
            while (tg_y != 0) {
                uint32_t bit = hi_bit(tg_y);
                tg_y &= ~(1u << bit);
                const bool two = tg_y != 0;
                uint32_t bit2 = two ? hi_bit(tg_y) : bit;      // odd count: the same record again (t2 == t1 never wins)
                tg_y &= ~(1u << bit2);
                uint32_t ti = tg_x + bit, ti2 = tg_x + bit2;
                const f32x4* tp = s.tris + (size_t)ti * kTriQuads;
                const f32x4* tp2 = s.tris + (size_t)ti2 * kTriQuads;
                f32x4 a = ldq(tp + 0), b = ldq(tp + 1), c = ldq(tp + 2);
                f32x4 a2 = ldq(tp2 + 0), b2 = ldq(tp2 + 1), c2 = ldq(tp2 + 2);
                float t1 = tr.closest, t2 = tr.closest;
                const bool h1 = intersect_tri(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), o, d, t_min, t1);
                const bool h2 = intersect_tri(mk3(a2.x, a2.y, a2.z), mk3(b2.x, b2.y, b2.z), mk3(c2.x, c2.y, c2.z), o, d, t_min, t2);
                if (h1) { tr.closest = t1; tr.best = ti; tr.bestw = f2u(a.w); best_inst = inst; }
                if (h2 && t2 < tr.closest) { tr.closest = t2; tr.best = ti2; tr.bestw = f2u(a2.w); best_inst = inst; }
            }
        }
        for (;;) {
            if (tr.ng_y > 0x00FFFFFFu) return false;    // a node group of the current level: visit it next step
            if (tr.ng_y != 0u) {                        // an instance list (top level only): enter its nearest instance
                uint32_t bit = hi_bit(tr.ng_y);
                tr.ng_y &= ~(1u << bit);
                stk.x[tr.sp] = tr.ng_x; stk.y[tr.sp] = tr.ng_y; tr.sp++;          // marker (+ the instances still to do)
                inst = f2u(ldq(s.tris + (size_t)(tr.ng_x + bit) * kTriQuads).w);
                // world -> object matrix and the BLAS root with four 128-bit loads (the record is 128-byte aligned; read member by
                // member this block was 87 instructions with 13 scalar loads, 12.7 % of the kernel on C5)
                const f32x4* ip = reinterpret_cast<const f32x4*>(s.inst + inst);
                const f32x4 i0 = ldq(ip + 3), i1 = ldq(ip + 4), i2 = ldq(ip + 5), i3 = ldq(ip + 6);
                const float iv[12] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w, i2.x, i2.y, i2.z, i2.w};
                o = xfm_pt(iv, ow); d = xfm_vec(iv, dw);
                set_dir(d);
                tr.ng_x = f2u(i3.x); tr.ng_y = 0x80000000u;                         // the BLAS root as a one-child group
                return false;
            }
            if (tr.sp == 0) return true;
            tr.sp--;
            tr.ng_x = stk.x[tr.sp]; tr.ng_y = stk.y[tr.sp];
            if (tr.ng_y <= 0x00FFFFFFu && inst != kMiss) {                          // marker: this BLAS is finished
                inst = kMiss; o = ow; d = dw;
                set_dir(d);
            }
        }
    }
};

// World-space shading inputs of an instanced hit: the flattened primitive id, the stored normal as loadPBRT's flattening would have
// computed it (vertex normal: unit(L n), utils/pbrt_loader.h:71-80; else the geometric normal of the transformed triangle) and the
// material index.  prim_shade[shade_base + k] = { n_obj.xyz | NaN when the mesh has no normals, material }.
PTX_HD void resolve_instanced_hit(const SceneView& s, const InstTraversal& it, uint32_t& prim, vec3& n, uint32_t& mat) {
    const f32x4* ip = reinterpret_cast<const f32x4*>(s.inst + it.best_inst);
    const f32x4 m0 = ldq(ip + 0), m1 = ldq(ip + 1), m2 = ldq(ip + 2), i3 = ldq(ip + 6);      // m[12] | root, prim_base, shade_base, flags
    const float m[12] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
    prim = f2u(i3.y) + it.tr.bestw;
    f32x4 ps = ldq(s.prim_shade + f2u(i3.z) + it.tr.bestw);
    mat = f2u(ps.w);
    if (ps.x == ps.x) n = unit(xfm_vec(m, mk3(ps.x, ps.y, ps.z)));
    else {
        const f32x4* tp = s.tris + (size_t)it.tr.best * kTriQuads;
        f32x4 b = ldq(tp + 1), c = ldq(tp + 2);
        n = unit(cross(xfm_vec(m, mk3(b.x, b.y, b.z)), xfm_vec(m, mk3(c.x, c.y, c.z))));
    }
}

template <int STACK>
PTX_HD uint32_t trace_closest_instanced(const SceneView& s, vec3 o, vec3 d, float t_min, float t_max, float& t_hit, uint32_t& prim) {
    InstTraversal it;
    TravStack<STACK> stk;
    it.begin(o, d, t_max);
    while (!it.template step<STACK>(s, stk, o, d, t_min)) {}
    t_hit = it.tr.closest;
    prim = kMiss;
    if (it.tr.best != kMiss) prim = s.inst[it.best_inst].prim_base + it.tr.bestw;
    return it.tr.best;
}

// Any hit along (o, d) in [t_min, t_max] through the two-level structure (Scene::occluded; trace_any in traverse.h has the end-point
// rule: a triangle accepts t == t_max, the halves of a quad do not -- quad_bits is indexed by the flattened primitive id).
template <int STACK>
PTX_HD bool trace_any_instanced(const SceneView& s, vec3 o, vec3 d, float t_min, float t_max) {
    if (!(t_max >= t_min) || !(t_max > 0.0f)) return false;
    const float bound = u2f(f2u(t_max) + 1u);
    InstTraversal it;
    TravStack<STACK> stk;
    it.begin(o, d, bound);
    for (;;) {
        const bool done = it.template step<STACK>(s, stk, o, d, t_min);
        if (it.tr.best != kMiss) {                         // a record accepted a hit below `bound`
            if (it.tr.closest < t_max || s.quad_bits == nullptr) return true;
            const uint32_t prim = f2u(ldq(reinterpret_cast<const f32x4*>(s.inst + it.best_inst) + 6).y) + it.tr.bestw;
            if (!((s.quad_bits[prim >> 5] >> (prim & 31u)) & 1u)) return true;
            it.tr.best = kMiss; it.tr.closest = bound;     // a quad half exactly at t_max does not count: keep looking
        }
        if (done) return false;
    }
}

}  // namespace ptx
