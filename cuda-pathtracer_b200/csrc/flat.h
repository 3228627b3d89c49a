// flat.h -- closest hit for TINY scenes (at most kFlatUnits planar units, e.g. the 16 quads of the Cornell box) without a
// tree: every ray runs one straight-line pass over all units, then ONE exact test.
//
// Replaces, for such scenes, Scene::intersect -> intersect_bvh_optimized (rendering/scene.h:39-110) exactly like the BVH8
// traversal of traverse.h does: the accepted hit is decided by the very same intersect_tri() arithmetic
// (Triangle::intersect / Quad::intersect, triangle.h:64-96, quad.h:49-132), so distances and primitive ids are bit-identical
// to the BVH8 path (exact ties between two different primitives excepted, as everywhere else).  Only the CULLING differs.
//
// Why: on a 4-node tree the per-lane state machine of k_path_persist spends 28 % of its issue slots in a triangle loop that
// runs 5.1 rounds per warp iteration for 1.9 tests per lane (lane efficiency 0.38) and a ray needs 1.5 iterations, so only
// 21 of 32 lanes finish a segment per iteration and shading runs at 0.35 (profiles/r1ze_c2_*).  Here the per-unit filter is
// uniform work (all lanes, unit data as constant-bank operands straight from the kernel parameters, no loads, no stack), every
// lane finishes its segment in the iteration it started it, and exactly one unit (two records) is tested exactly.
//
// A unit = one planar primitive: a reference quad (both halves), or two coplanar triangles sharing an edge, or a single
// triangle.  Per unit the filter intersects the ray with the unit's plane and checks the hit point against the unit's
// bounding rectangle IN the plane (exact fit for rectangles / parallelograms' bounding box), with a margin that grows as
// the ray grazes the plane:      m = m0 + m1 / |n.d|.
// The filter is CONSERVATIVE: whatever intersect_tri() accepts lies inside the candidate set (the margins cover the
// rounding of both computations, see flat_margins()).  tlo = t - m is then a lower bound of the unit's exact distance; the
// candidate with the smallest tlo is tested exactly, and if its exact distance is below the second smallest tlo nothing
// else can be closer.  Otherwise (two candidates within their margins: rays into an edge between two walls; a grazing
// plane; a false candidate in front) the lane tests every candidate exactly -- rare (measured on the host emulation and
// reported by the GPU tests).
#pragma once
#include "layout.h"
#include "traverse.h"

namespace ptx {

constexpr int kFlatUnits = 20;      // the Cornell box: 15 planar quads + the skewed left wall as two triangles = 17
constexpr int kFlatRecQuads = 6;       // two 48-byte edge records per unit (a single triangle is followed by a null record)

struct FlatUnit {                      // 64 bytes, read as constant-bank operands
    float nx, ny, nz, c;               // unit normal, plane offset (n.x = c)
    float cx, cy, cz, h1;              // centre of the bounding rectangle (a point of the plane), half extent along a1
    float a1x, a1y, a1z, h2;           // in-plane unit axes a1, a2 = n x a1; half extent along a2
    float a2x, a2y, a2z, pad;
};

// Two units side by side, every constant as a (unit 2j, unit 2j + 1) pair: the device filter handles a pair per step with
// packed FP32x2 arithmetic (FFMA2 takes the constant pair from a uniform-register pair and the ray component as a
// broadcast scalar), which halves the issue slots of the dot products.  mc = -c.
struct FlatPair {                      // 128 bytes
    float nx[2], ny[2], nz[2], mc[2];
    float ncx[2], ncy[2], ncz[2], h1[2];     // MINUS the centre of the bounding rectangle
    float a1x[2], a1y[2], a1z[2], h2[2];
    float a2x[2], a2y[2], a2z[2], pad[2];
};

struct FlatScene {
    FlatPair p[kFlatUnits / 2];
    const f32x4* recs;                 // kFlatRecQuads float4 per unit: { v0, prim }, { e1 }, { e2 } twice
    uint32_t n_units;
    float m0, m1;                      // margins, see flat_margins()
    uint32_t key_mask;                 // 0xFFFFFFE0: clears the 5 key bits that carry the unit index (a register operand of LOP3)
};

// unit k of the list; slots never set stay all-zero: n = 0 -> 1 / 0 = inf -> NaN / inf everywhere, so the loop bound
// n_units (not the data) decides what is evaluated
inline void flat_set_unit(FlatScene& fs, int k, const FlatUnit& u) {
    FlatPair& p = fs.p[k >> 1];
    const int i = k & 1;
    p.nx[i] = u.nx; p.ny[i] = u.ny; p.nz[i] = u.nz; p.mc[i] = -u.c;
    p.ncx[i] = -u.cx; p.ncy[i] = -u.cy; p.ncz[i] = -u.cz; p.h1[i] = u.h1;
    p.a1x[i] = u.a1x; p.a1y[i] = u.a1y; p.a1z[i] = u.a1z; p.h2[i] = u.h2;
    p.a2x[i] = u.a2x; p.a2y[i] = u.a2y; p.a2z[i] = u.a2z; p.pad[i] = 0.f;
}

// Margins for rays whose origins and the scene lie within `scale` of the coordinate origin.
//   error of t = num / den:  num = c - n.o cancels to ~ 4 ulp of `scale`, den to ~ 3 ulp of 1, |t| <= 2 * scale:
//                            |dt| <= (2.4e-7 + 2 * 1.8e-7 + rcp.approx 1.2e-7 * 2) * scale / |den|  < 1e-6 * scale / |den|
//   tolerance of intersect_tri at its borders (u, v computed as f * dot(s, h) with f = 1 / (e1 . (d x e2))): a point it accepts
//   lies within ~ 3e-7 * (|s| + |e|) / (sin(edge angle) * |den|) of the primitive:                           < 3e-6 * scale / |den|
//   storage rounding of the unit constants (host double -> float) and of the in-plane projections:         < 1e-6 * scale
// m0 = 1e-5 * scale and m1 = 1e-5 * scale leave a factor > 2 on top of their sum.
inline void flat_margins(float scale, float& m0, float& m1) { m0 = 1e-5f * scale; m1 = 1e-5f * scale; }
constexpr uint32_t kFlatKeyMask = 0xFFFFFFE0u;

// 1 / x for the filter: one MUFU.RCP, denormals flushed (a denormal n.d becomes 0 -> 1 / 0 = inf -> infinite margin -> candidate).
// rcp.approx WITHOUT .ftz costs two range tests, two selects and two scalings around the MUFU.
PTX_HD float rcp_ftz(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.0f / x;
#endif
}

// exact test of unit k (both records, the second replaces the first only if strictly closer: quad.h:76, 106)
PTX_HD void flat_test_unit(const f32x4* recs, uint32_t k, vec3 o, vec3 d, float t_min, float& closest, uint32_t& bestw) {
    const f32x4* tp = recs + (size_t)k * kFlatRecQuads;
    f32x4 a = ldq(tp + 0), b = ldq(tp + 1), c = ldq(tp + 2);
    f32x4 a2 = ldq(tp + 3), b2 = ldq(tp + 4), c2 = ldq(tp + 5);
    if (intersect_tri(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), o, d, t_min, closest)) bestw = f2u(a.w);
    if (intersect_tri(mk3(a2.x, a2.y, a2.z), mk3(b2.x, b2.y, b2.z), mk3(c2.x, c2.y, c2.z), o, d, t_min, closest)) bestw = f2u(a2.w);
}

struct FlatCounters {
    uint32_t rays, slow, unit_tests;
};

// packed pair helpers: FFMA2 / FMUL2 / FADD2 on the device (sm_100: one issue slot for two IEEE operations), the same two
// operations one after the other on the host
struct f2 {
    float x, y;
};
PTX_HD f2 mk2(float a, float b) { f2 r; r.x = a; r.y = b; return r; }
PTX_HD f2 ld2(const float* p) {
#if defined(__CUDA_ARCH__)
    const float2 v = *reinterpret_cast<const float2*>(p);
    return mk2(v.x, v.y);
#else
    return mk2(p[0], p[1]);
#endif
}
PTX_HD f2 fma2(f2 a, f2 b, f2 c) {
#if defined(__CUDA_ARCH__)
    const float2 r = __ffma2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y), make_float2(c.x, c.y));
    return mk2(r.x, r.y);
#else
    return mk2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
#endif
}
PTX_HD f2 mul2(f2 a, f2 b) {
#if defined(__CUDA_ARCH__)
    const float2 r = __fmul2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
    return mk2(r.x, r.y);
#else
    return mk2(a.x * b.x, a.y * b.y);
#endif
}
PTX_HD f2 add2(f2 a, f2 b) {
#if defined(__CUDA_ARCH__)
    const float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
    return mk2(r.x, r.y);
#else
    return mk2(a.x + b.x, a.y + b.y);
#endif
}

// min / max that drop a NaN SECOND operand and never produce one from a non-NaN first operand: what FMNMX does on the device.
// (glibc's fminf returns NaN for a signalling NaN, and (-inf & ~31) | k is one.)
PTX_HD float min_nn(float a, float b) {
#if defined(__CUDA_ARCH__)
    return fminf(a, b);
#else
    return b < a ? b : a;
#endif
}
PTX_HD float max_nn(float a, float b) {
#if defined(__CUDA_ARCH__)
    return fmaxf(a, b);
#else
    return b > a ? b : a;
#endif
}

// one unit's share of the selection: key = lower bound t - m with the unit index in its low 5 bits, so ONE min tracks both
// the smallest bound and whose it is; t2 = second smallest key.
PTX_HD void flat_select(int k, float tpm, float tmm, float e1, float e2, float hm1, float hm2, float t_min, uint32_t key_mask, float& tbest,
                        float& t2) {
    // tpm = t + m, tmm = t - m.  A NaN (n.d = 0 after flushing: t = 0 * inf, or the all-zero unit that pads an odd list) makes
    // the unit a NON-candidate: it cannot be hit, intersect_tri rejects |det| = 2 area |n.d| < 1e-8.  A grazing but finite
    // 1 / |n.d| just gives a large finite margin.
    const bool cand = (fabsf(e1) <= hm1) && (fabsf(e2) <= hm2) && (tpm >= t_min);
#if defined(__CUDA_ARCH__)
    uint32_t kb;      // (bits & mask) | k as ONE LOP3: the mask sits in a register (it comes from the parameter block), k is the immediate
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(kb) : "r"(f2u(tmm)), "r"(key_mask), "r"((uint32_t)k));
    const float key = u2f(kb);
#else
    const float key = u2f((f2u(tmm) & key_mask) | (uint32_t)k);
#endif
    if (cand) {
        t2 = min_nn(t2, max_nn(tbest, key));
        tbest = min_nn(tbest, key);
    }
}

// Closest hit of (o, d) in [t_min, FLT_MAX).  closest / bestw (primitive id or kMiss) as Traversal leaves them.
template <bool COUNT>
PTX_HD void flat_closest(const FlatScene& fs, vec3 o, vec3 d, float t_min, float& closest, uint32_t& bestw, FlatCounters* ctr) {
    const float kInf = u2f(0x7F800000u);
    float tbest = kInf, t2 = kInf;
    const uint32_t n = fs.n_units;
    const f2 ox = mk2(o.x, o.x), oy = mk2(o.y, o.y), oz = mk2(o.z, o.z);
    const f2 dx = mk2(d.x, d.x), dy = mk2(d.y, d.y), dz = mk2(d.z, d.z);
    const uint32_t key_mask = fs.key_mask;                            // 0xFFFFFFE0
    const f2 m1 = mk2(fs.m1, fs.m1), m0 = mk2(fs.m0, fs.m0), minus1 = mk2(-1.0f, -1.0f);
#pragma unroll
    for (int j = 0; j < kFlatUnits / 2; j++) {
        if ((uint32_t)(2 * j) < n) {                                  // uniform; an odd list ends with an all-zero unit (NaN key: dropped)
            const FlatPair& u = fs.p[j];
            const f2 den = fma2(ld2(u.nz), dz, fma2(ld2(u.ny), dy, mul2(ld2(u.nx), dx)));
            const f2 numn = fma2(ld2(u.nz), oz, fma2(ld2(u.ny), oy, fma2(ld2(u.nx), ox, ld2(u.mc))));     // n.o - c
            const f2 rn = mk2(rcp_ftz(-den.x), rcp_ftz(-den.y));
            const f2 t = mul2(numn, rn);
            const f2 m = fma2(mk2(fabsf(rn.x), fabsf(rn.y)), m1, m0);                                      // m0 + m1 / |n.d|
            const f2 px = add2(fma2(t, dx, ox), ld2(u.ncx)), py = add2(fma2(t, dy, oy), ld2(u.ncy)), pz = add2(fma2(t, dz, oz), ld2(u.ncz));
            const f2 e1 = fma2(ld2(u.a1z), pz, fma2(ld2(u.a1y), py, mul2(ld2(u.a1x), px)));
            const f2 e2 = fma2(ld2(u.a2z), pz, fma2(ld2(u.a2y), py, mul2(ld2(u.a2x), px)));
            const f2 hm1 = add2(ld2(u.h1), m), hm2 = add2(ld2(u.h2), m);
            const f2 tpm = add2(t, m), tmm = fma2(m, minus1, t);
            flat_select(2 * j, tpm.x, tmm.x, e1.x, e2.x, hm1.x, hm2.x, t_min, key_mask, tbest, t2);
            flat_select(2 * j + 1, tpm.y, tmm.y, e1.y, e2.y, hm1.y, hm2.y, t_min, key_mask, tbest, t2);
        }
    }
    closest = FLT_MAX; bestw = kMiss;
    if (COUNT) ctr->rays++;
    if (tbest < kInf) {
        const uint32_t kbest = f2u(tbest) & 31u;
        flat_test_unit(fs.recs, kbest, o, d, t_min, closest, bestw);
        if (COUNT) ctr->unit_tests++;
        // the keys carry the unit index in their low bits, i.e. they may exceed their bound by < 32 ulp: take 2^-18 off
        if (!(bestw != kMiss && closest < fmaf(t2, -3.8146973e-06f, t2)) && t2 < kInf) {
            // not provably the closest (two candidates within their margins: a ray into the edge between two walls; a grazing
            // plane; a false candidate in front): every other unit gets its exact test.  Rare (< 1e-3 of the rays).
            if (COUNT) ctr->slow++;
            for (uint32_t k = 0; k < n; k++) {
                if (k == kbest) continue;
                flat_test_unit(fs.recs, k, o, d, t_min, closest, bestw);
                if (COUNT) ctr->unit_tests++;
            }
        }
    }
}

}  // namespace ptx
