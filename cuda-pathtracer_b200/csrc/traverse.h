// traverse.h -- closest-hit traversal of the compressed BVH8 and the triangle test.
//
// Replaces Scene::intersect -> intersect_bvh_optimized (rendering/scene.h:39-110) and
// Triangle::intersect / Quad::intersect (rendering/triangle.h:64-96, rendering/quad.h:49-132).
// Acceptance rule kept from the reference: a candidate replaces the current hit iff
// t > 1e-8, t >= t_min and t < closest (strict, scene.h:89-90); |det| < 1e-8 rejects; u in [0,1],
// v >= 0, u+v <= 1 inclusive (triangle.h:71-80).
//
// Written __host__ __device__ so tests can compile the very same source for the CPU and compare
// it with the oracle without a GPU; the product only ever calls it from CUDA kernels.
#pragma once
#include "layout.h"

namespace ptx {

PTX_HD uint32_t hi_bit(uint32_t x) {   // index of the highest set bit, x != 0
#if defined(__CUDA_ARCH__)
    return 31u - (uint32_t)__clz((int)x);
#else
    return 31u - (uint32_t)__builtin_clz(x);
#endif
}
PTX_HD uint32_t popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return (uint32_t)__popc(x);
#else
    return (uint32_t)__builtin_popcount(x);
#endif
}
PTX_HD uint32_t byte_of(uint32_t w, int i) { return (w >> (8 * i)) & 0xFFu; }
// byte i of w as a float.  Device: one PRMT builds 0x4B0000bb = 2^23 + b, one FADD removes 2^23 (exact),
// which keeps the conversion off the quarter-rate XU pipe that I2F.U8 runs on (ncu r1a: XU 60-67 % busy).
PTX_HD float byte_to_float(uint32_t w, int i, uint32_t magic = 0x4B000000u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__byte_perm(w, magic, 0x7440u + (uint32_t)i)) - 8388608.0f;
#else
    (void)magic;
    return (float)byte_of(w, i);
#endif
}
// per byte: 0x80 -> 0xFF, 0x00 -> 0x00
PTX_HD uint32_t sign_extend_s8x4(uint32_t x) { return ((x >> 7) & 0x01010101u) * 0xFFu; }

struct TraceCounters {
    uint32_t nodes, tris;
};

// 1/x for the slab tests only (never for hit distances): one MUFU.RCP on the device (<= 1 ulp) instead of
// the 16-instruction IEEE division; the child boxes are quantised outwards by up to a whole cell, so the
// last ulp of 1/d does not decide a box test.
PTX_HD float rcp_fast(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.0f / x;
#endif
}

// Moeller-Trumbore on a precomputed edge record.  Returns true and updates (closest, u, v) on accept.
PTX_HD bool intersect_tri(vec3 v0, vec3 e1, vec3 e2, vec3 o, vec3 d, float t_min, float& closest) {
    const float EPSILON = 1e-8f;
    vec3 h = cross(d, e2);
    float a = dot(e1, h);
#if defined(__CUDA_ARCH__)
    // Straight-line form for the GPU: the same operations in the same order as the early-out form below (so every
    // accepted t is bit-identical), with the rejections collected into one predicate.  With ~12 of 32 lanes active
    // an early exit almost never frees the warp, while its three nested divergent branches cost BSSY/BSYNC pairs
    // and keep ptxas from overlapping the two cross products (ncu r1i: branch_resolving 0.66 per issue).
    float f = rcp_exact(a);
    vec3 s = o - v0;
    float u = f * dot(s, h);
    vec3 q = cross(s, e1);
    float v = f * dot(d, q);
    float t = f * dot(e2, q);
    bool ok = !(fabsf(a) < EPSILON) && !(u < 0.0f || u > 1.0f) && !(v < 0.0f || u + v > 1.0f) && (t > EPSILON && t >= t_min && t < closest);
    if (ok) closest = t;
    return ok;
#else
    if (fabsf(a) < EPSILON) return false;
    float f = 1.0f / a;
    vec3 s = o - v0;
    float u = f * dot(s, h);
    if (u < 0.0f || u > 1.0f) return false;
    vec3 q = cross(s, e1);
    float v = f * dot(d, q);
    if (v < 0.0f || u + v > 1.0f) return false;
    float t = f * dot(e2, q);
    if (t > EPSILON && t >= t_min && t < closest) { closest = t; return true; }
    return false;
#endif
}

// Resumable closest-hit traversal.  One step() = visit ONE wide node (the nearest pending child of the
// current node group), test the triangles of the leaves it hit, then fetch the next node group from
// the stack.  Kernels that keep a path per lane call step() once per loop iteration and shade as soon
// as a lane's traversal finishes, so a lane whose ray needed one node visit does not idle while its
// neighbours visit three (k_path_loop).  trace_closest() below just runs step() to completion.
// STACK entries of 8 bytes live in registers/local memory.
// The stack lives in its own object so that the scalar members of Traversal stay in registers
// (one struct holding both was kept in local memory as a whole: STL/LDL of idx.. in ncu r1c).
template <int STACK>
struct TravStack {
    uint32_t x[STACK], y[STACK];
};

struct Traversal {
    int sp;
    float closest;
    uint32_t best;                 // triangle slot (index into tris) of the closest accepted hit, or kMiss
    uint32_t bestw;                // its primitive id (word 3 of the record), kept so that shading need not fetch the record again
    uint32_t ng_x, ng_y;           // node group: base index, (hit bits 31..24 | imask 7..0)
    float idx, idy, idz;           // 1/d with zero components replaced by +-2^-80

    PTX_HD void begin(vec3 d, float t_max) {
        sp = 0; closest = t_max; best = kMiss; bestw = kMiss;
        const float ooeps = 8.271806e-25f;   // 2^-80: keeps 1/d finite for axis-parallel rays
        // The sign of the substitute follows `d < 0`, the very test visit() orders the slab planes by: copysign() made 1/d
        // negative for d = -0.0f while visit() took the component for positive, so entry and exit plane were swapped and a
        // ray with a negative-zero component missed everything (1 segment in 1.5e7 on the Cornell box; found by comparing
        // with flat_closest, tests/test_host_logic.py::test_negative_zero_direction).
        idx = rcp_fast(fabsf(d.x) > ooeps ? d.x : (d.x < 0.0f ? -ooeps : ooeps));
        idy = rcp_fast(fabsf(d.y) > ooeps ? d.y : (d.y < 0.0f ? -ooeps : ooeps));
        idz = rcp_fast(fabsf(d.z) > ooeps ? d.z : (d.z < 0.0f ? -ooeps : ooeps));
        ng_x = 0; ng_y = 0x80000000u;        // the root as a one-child group
    }

    // Visit ONE wide node: the nearest pending inner child of the current node group.  Leaves the hit inner children in
    // (ng_x, ng_y) -- pushing the rest of the old group first -- and returns the hit triangles as (tg_x, tg_y): base slot
    // and one bit per triangle, nearest first from the top.  step() below and the CTA-wavefront kernel (k_path_cta, which
    // hands the triangle tests to other lanes) share it.
    template <bool COUNT, class STK>
    PTX_HD void visit(const SceneView& s, STK& stk, vec3 o, vec3 d, float t_min, uint32_t& tg_x_out, uint32_t& tg_y_out, TraceCounters* ctr) {
        // octant: bit set where the direction is negative; slot priority = slot ^ octinv
        uint32_t oct = (d.x < 0.0f ? 4u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 1u : 0u);
        uint32_t octinv = 7u - oct;
        uint32_t octinv4 = octinv * 0x01010101u;
        uint32_t tg_x = 0, tg_y = 0;         // triangle group: base index, hit bits 23..0
        // 0x3F800000 comes from the kernel parameter block (a constant-bank operand of PRMT): with the constant known at
        // compile time ptxas spends PRMT's single immediate on it and materialises every byte selector with a MOV
        const uint32_t magic = s.byte_magic;
        (void)magic;

        {   // on entry the node group always holds a pending inner child (begin() / the pop below guarantee it)
            uint32_t imask = ng_y & 0xFFu;
            uint32_t bit = hi_bit(ng_y);
            ng_y &= ~(1u << bit);
            if (ng_y > 0x00FFFFFFu) { stk.x[sp] = ng_x; stk.y[sp] = ng_y; sp++; }
            uint32_t slot = (bit - 24u) ^ octinv;
            uint32_t rel = popc32(imask & ~(0xFFFFFFFFu << slot));
            const u32x4* np = s.nodes + (size_t)(ng_x + rel) * kNodeQuads;
            u32x4 n0 = ldq(np + 0), n1 = ldq(np + 1), n2 = ldq(np + 2), n3 = ldq(np + 3), n4 = ldq(np + 4);
            if (COUNT) ctr->nodes++;

            float px = u2f(n0.x), py = u2f(n0.y), pz = u2f(n0.z);
            uint32_t ex = n0.w & 0xFFu, ey = (n0.w >> 8) & 0xFFu, ez = (n0.w >> 16) & 0xFFu;
            ng_x = n1.x; tg_x = n1.y;
            float sx = u2f(ex << 23) * idx, sy = u2f(ey << 23) * idy, sz = u2f(ez << 23) * idz;
            float ox = (px - o.x) * idx, oy = (py - o.y) * idy, oz = (pz - o.z) * idz;

            uint32_t hitmask = 0;
#if defined(__CUDA_ARCH__)
            // Device form of the slab tests.  A quantised byte q becomes the float v = 1 + q 2^-15 with ONE PRMT
            // (0x3F80qq00), and the "- 1" is folded into the additive term:  t = q s + o = v S + B  with S = 2^15 s,
            // B = o - S.  B is rounded towards -inf for entry planes and +inf for exit planes, so the test stays
            // conservative (the rounding is <= 2^-8 of one quantisation cell).  Entry and exit plane of an axis go
            // through one packed FFMA2 (fma.rn.f32x2, sm_100): 3 issue slots per child instead of 6 FFMA + 6 FADD.
            const float Sx = sx * 32768.0f, Sy = sy * 32768.0f, Sz = sz * 32768.0f;
            const float2 S2x = make_float2(Sx, Sx), S2y = make_float2(Sy, Sy), S2z = make_float2(Sz, Sz);
            const float2 B2x = make_float2(__fsub_rd(ox, Sx), __fsub_ru(ox, Sx));
            const float2 B2y = make_float2(__fsub_rd(oy, Sy), __fsub_ru(oy, Sy));
            const float2 B2z = make_float2(__fsub_rd(oz, Sz), __fsub_ru(oz, Sz));
#pragma unroll
            for (int half = 0; half < 2; half++) {
                uint32_t meta4 = half ? n1.w : n1.z;
                if (meta4 == 0u) continue;                 // four empty slots (small nodes fill the low slots first); dropping the
                                                           // branch was measured neutral (C2 18.88 -> 18.90 G, 8 bytes of spills)
                uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
                uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
                uint32_t m4 = meta4 ^ (octinv4 & inner_mask4 & 0x07070707u);   // per byte: bits 7..5 unary count, 4..0 bit index
                uint32_t qlox = half ? n2.y : n2.x, qloy = half ? n2.w : n2.z, qloz = half ? n3.y : n3.x;
                uint32_t qhix = half ? n3.w : n3.z, qhiy = half ? n4.y : n4.x, qhiz = half ? n4.w : n4.z;
                uint32_t nearx = d.x < 0.0f ? qhix : qlox, farx = d.x < 0.0f ? qlox : qhix;
                uint32_t neary = d.y < 0.0f ? qhiy : qloy, fary = d.y < 0.0f ? qloy : qhiy;
                uint32_t nearz = d.z < 0.0f ? qhiz : qloz, farz = d.z < 0.0f ? qloz : qhiz;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t sel = 0x7604u + ((uint32_t)j << 4);       // {magic.3, magic.2, w.j, magic.0}
                    float2 tx = __ffma2_rn(make_float2(__uint_as_float(__byte_perm(nearx, magic, sel)), __uint_as_float(__byte_perm(farx, magic, sel))), S2x, B2x);
                    float2 ty = __ffma2_rn(make_float2(__uint_as_float(__byte_perm(neary, magic, sel)), __uint_as_float(__byte_perm(fary, magic, sel))), S2y, B2y);
                    float2 tz = __ffma2_rn(make_float2(__uint_as_float(__byte_perm(nearz, magic, sel)), __uint_as_float(__byte_perm(farz, magic, sel))), S2z, B2z);
                    float cmin = fmaxf(fmaxf(tx.x, ty.x), fmaxf(tz.x, t_min));
                    float cmax = fminf(fminf(tx.y, ty.y), fminf(tz.y, closest));
                    uint32_t mj = __byte_perm(m4, 0u, 0x4440u + (uint32_t)j);
                    uint32_t bits;
                    asm("shf.l.wrap.b32 %0, 0, %1, %2;" : "=r"(bits) : "r"(mj >> 5), "r"(mj));   // (mj >> 5) << (mj & 31)
                    if (cmin <= cmax) hitmask |= bits;
                }
            }
#else
            for (int half = 0; half < 2; half++) {
                uint32_t meta4 = half ? n1.w : n1.z;
                if (meta4 == 0u) continue;
                uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
                uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
                uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
                uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
                uint32_t qlox = half ? n2.y : n2.x, qloy = half ? n2.w : n2.z, qloz = half ? n3.y : n3.x;
                uint32_t qhix = half ? n3.w : n3.z, qhiy = half ? n4.y : n4.x, qhiz = half ? n4.w : n4.z;
                uint32_t nearx = d.x < 0.0f ? qhix : qlox, farx = d.x < 0.0f ? qlox : qhix;
                uint32_t neary = d.y < 0.0f ? qhiy : qloy, fary = d.y < 0.0f ? qloy : qhiy;
                uint32_t nearz = d.z < 0.0f ? qhiz : qloz, farz = d.z < 0.0f ? qloz : qhiz;
                for (int j = 0; j < 4; j++) {
                    float tminx = (float)byte_of(nearx, j) * sx + ox, tmaxx = (float)byte_of(farx, j) * sx + ox;
                    float tminy = (float)byte_of(neary, j) * sy + oy, tmaxy = (float)byte_of(fary, j) * sy + oy;
                    float tminz = (float)byte_of(nearz, j) * sz + oz, tmaxz = (float)byte_of(farz, j) * sz + oz;
                    float cmin = fmaxf(fmaxf(tminx, tminy), fmaxf(tminz, t_min));
                    float cmax = fminf(fminf(tmaxx, tmaxy), fminf(tmaxz, closest));
                    if (cmin <= cmax) hitmask |= byte_of(child_bits4, j) << byte_of(bit_index4, j);
                }
            }
#endif
            ng_y = (hitmask & 0xFF000000u) | (n0.w >> 24);
            tg_y = hitmask & 0x00FFFFFFu;
            // (Software prefetch of the hit triangles / the nearest child with prefetch.global.L1 or .L2 was measured on C3
            //  and HALVED throughput -- 3.9 -> 1.7 / 1.5 G segments/s -- so the records are simply loaded when needed.)
        }
        tg_x_out = tg_x; tg_y_out = tg_y;
    }

    // Fetch the next node group from the stack when the current one holds no inner child any more.
    // Returns true when the traversal is complete (nothing left to visit).
    template <class STK>
    PTX_HD bool next_group(STK& stk) {
        if (ng_y <= 0x00FFFFFFu) {
            if (sp == 0) return true;
            sp--;
            ng_x = stk.x[sp]; ng_y = stk.y[sp];
        }
        return false;
    }

    // returns true when the traversal is complete (closest / best are final)
    // PAIR: the triangle loop takes two records per round (both tested against the distance the round started with, then
    // committed in order -- the same accepted hits as one at a time, bit for bit).  Two independent fetches in flight per
    // lane are worth +6 % where the records come from L2 / HBM (C3 7.13 -> 7.56 G) and +2 % at 41 K triangles; on the
    // 32-triangle Cornell box the extra work of odd counts costs 0.6 % (C2), so the launcher picks it by scene size.
    template <bool COUNT, int STACK, bool PAIR = false>
    PTX_HD bool step(const SceneView& s, TravStack<STACK>& stk, vec3 o, vec3 d, float t_min, TraceCounters* ctr) {
        uint32_t tg_x, tg_y;
        visit<COUNT>(s, stk, o, d, t_min, tg_x, tg_y, ctr);
        // (A software-pipelined form of this loop -- the record of test k+1 requested before test k is evaluated, 12 more live
        //  registers, still 80 in total -- was measured: C2 17.1 -> 16.1 G, C3 6.58 -> 6.35 G.  The resident warps already
        //  cover the fetch; the extra moves and the longer loop body only cost issue slots.)
        if constexpr (PAIR) {
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
                if (COUNT) ctr->tris += two ? 2u : 1u;
                float t1 = closest, t2 = closest;
                const bool h1 = intersect_tri(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), o, d, t_min, t1);
                const bool h2 = intersect_tri(mk3(a2.x, a2.y, a2.z), mk3(b2.x, b2.y, b2.z), mk3(c2.x, c2.y, c2.z), o, d, t_min, t2);
                if (h1) { closest = t1; best = ti; bestw = f2u(a.w); }
                if (h2 && t2 < closest) { closest = t2; best = ti2; bestw = f2u(a2.w); }
            }
        } else
        while (tg_y != 0) {
            uint32_t bit = hi_bit(tg_y);
            tg_y &= ~(1u << bit);
            uint32_t ti = tg_x + bit;
            const f32x4* tp = s.tris + (size_t)ti * kTriQuads;
            f32x4 a = ldq(tp + 0), b = ldq(tp + 1), c = ldq(tp + 2);
            if (COUNT) ctr->tris++;
            if (intersect_tri(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), o, d, t_min, closest)) { best = ti; bestw = f2u(a.w); }
        }

        return next_group(stk);
        // (Prefetching the node the next step will visit -- prefetch.global.L2 or .L1 on its two cache lines, issued here --
        //  was measured too: C3 6.9 -> 6.2 G, C2 17.6 -> 15.1 G, the C5-shaped scene 1.26 -> 0.86 G.  Like the triangle
        //  prefetch above it only costs; the warps already in flight hide what latency can be hidden.
        //  Staging the next node in a per-lane shared-memory slot with cp.async + wait_group was bit-identical and also slower:
        //  C3 6.9 -> 6.6 G, C2 17.6 -> 16.5 G, C5-shaped 1.26 -> 1.19 G.)
    }
};

// Closest hit along (o, d) in [t_min, t_max).  Returns the triangle slot (index into tris) or kMiss;
// t_hit receives the distance.
template <int STACK, bool COUNT>
PTX_HD uint32_t trace_closest(const SceneView& s, vec3 o, vec3 d, float t_min, float t_max, float& t_hit,
                              TraceCounters* ctr) {
    Traversal tr;
    TravStack<STACK> stk;
    tr.begin(d, t_max);
    while (!tr.template step<COUNT, STACK>(s, stk, o, d, t_min, ctr)) {}
    t_hit = tr.closest;
    return tr.best;
}

// Any hit along (o, d) in [t_min, t_max]: the traversal stops at the first accepted record (Scene::occluded_bvh, scene.h:152-190).
// The answer does not depend on the visiting order.  End-point rule of the reference: Triangle::intersect accepts
// t_min <= t <= t_max (triangle.h:83), the two halves of a Quad only t < t_max (quad.h:76, 106: `t < closest_t` with
// closest_t = t_max).  So the records are tested against the next float above t_max, and a hit at exactly t_max counts
// unless its primitive is a quad (s.quad_bits: one bit per reference primitive, read only in that rare case).
template <int STACK>
PTX_HD bool trace_any(const SceneView& s, vec3 o, vec3 d, float t_min, float t_max) {
    if (!(t_max >= t_min) || !(t_max > 0.0f)) return false;
    const float bound = u2f(f2u(t_max) + 1u);          // t < bound  <=>  t <= t_max   (t_max > 0, finite)
    Traversal tr;
    TravStack<STACK> stk;
    tr.begin(d, bound);
    for (;;) {
        uint32_t tg_x, tg_y;
        tr.template visit<false>(s, stk, o, d, t_min, tg_x, tg_y, nullptr);
        while (tg_y != 0) {
            uint32_t bit = hi_bit(tg_y);
            tg_y &= ~(1u << bit);
            const f32x4* tp = s.tris + (size_t)(tg_x + bit) * kTriQuads;
            f32x4 a = ldq(tp + 0), b = ldq(tp + 1), c = ldq(tp + 2);
            float t = bound;
            if (intersect_tri(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), o, d, t_min, t)) {
                if (t < t_max) return true;
                const uint32_t prim = f2u(a.w);
                if (!(s.quad_bits != nullptr && ((s.quad_bits[prim >> 5] >> (prim & 31u)) & 1u))) return true;
            }
        }
        if (tr.next_group(stk)) return false;
    }
}

}  // namespace ptx
