// rng.h -- the two random-number modes of the integrator.
//
//  * XORWOW "compat" mode reproduces, bit for bit, the per-pixel streams the reference creates with
//    curand_init(2023 + pix, pix, 0, &state[pix]) (rendering/integrator.h:274-280) and consumes with
//    curand_uniform (integrator.h:63-64, 210, 384-385).  The generator is Marsaglia's xorwow with a
//    Weyl counter; `subsequence = pix` means a jump of pix * 2^67 draws, done here with our own
//    GF(2) jump matrices J[b] = T^(2^(67+b)) (T = one-step transition of the 160-bit state), built on
//    the host by repeated squaring at context creation and applied on the device bit by bit.
//    (cuRAND ships equivalent precomputed tables; none of its code or tables is used.)
//  * Philox4x32-10 "fast" mode: counter-based, keyed by (seed), counter = (path index, draw block),
//    so any GPU can render any sample range and get the same image.
//
// Both map a 32-bit draw x to a float in (0,1] as x * 2^-32 + 2^-33 (cuRAND's _curand_uniform).
#pragma once
#include "common.h"

namespace ptx {

struct Xorwow {
    uint32_t v0, v1, v2, v3, v4, d;
};

PTX_HD uint32_t xorwow_next(Xorwow& s) {
    uint32_t t = s.v0 ^ (s.v0 >> 2);
    s.v0 = s.v1; s.v1 = s.v2; s.v2 = s.v3; s.v3 = s.v4;
    s.v4 = (s.v4 ^ (s.v4 << 4)) ^ (t ^ (t << 1));
    s.d += 362437u;
    return s.v4 + s.d;
}
PTX_HD float u32_to_unit(uint32_t x) { return (float)x * 2.3283064e-10f + (2.3283064e-10f / 2.0f); }
PTX_HD float xorwow_uniform(Xorwow& s) { return u32_to_unit(xorwow_next(s)); }

// seed scrambling of curand_init for XORWOW (before the subsequence jump)
PTX_HD Xorwow xorwow_seed(uint64_t seed) {
    uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
    uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
    uint32_t t0 = 1099087573u * s0;
    uint32_t t1 = 2591861531u * s1;
    Xorwow s;
    s.d = 6615241u + t1 + t0;
    s.v0 = 123456789u + t0;
    s.v1 = 362436069u ^ t0;
    s.v2 = 521288629u + t1;
    s.v3 = 88675123u ^ t1;
    s.v4 = 5783321u + t0;
    return s;
}

constexpr int kXorwowRows = 160;     // state bits
constexpr int kXorwowWords = 5;
constexpr int kXorwowJumps = 64;     // J[b], b = 0..63: one per bit of the subsequence number
constexpr int kXorwowJumpWords = kXorwowRows * kXorwowWords;   // 800 words per matrix

// row-vector * matrix over GF(2): row r of `m` is the image of state bit r
PTX_HD void xorwow_apply(Xorwow& s, const uint32_t* m) {
    uint32_t in[5] = {s.v0, s.v1, s.v2, s.v3, s.v4};
    uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
    for (int w = 0; w < 5; w++) {
        uint32_t bits = in[w];
        const uint32_t* row = m + (size_t)w * 32 * 5;
        for (int j = 0; j < 32; j++, row += 5) {
            uint32_t mask = 0u - ((bits >> j) & 1u);
            r0 ^= row[0] & mask; r1 ^= row[1] & mask; r2 ^= row[2] & mask; r3 ^= row[3] & mask; r4 ^= row[4] & mask;
        }
    }
    s.v0 = r0; s.v1 = r1; s.v2 = r2; s.v3 = r3; s.v4 = r4;
}

// host: fill `out` (kXorwowJumps * kXorwowJumpWords words) -- see rng.cpp
void xorwow_build_jump_tables(uint32_t* out);
// host: `out` (kXorwowJumpWords words) = T^n, the transition of n draws (the Weyl counter advances by 362437 * n besides); rng.cpp
void xorwow_build_skip_matrix(uint64_t n, uint32_t* out);

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC11)
// ---------------------------------------------------------------------------------------------
struct Philox4 {
    uint32_t x, y, z, w;
};
PTX_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}
PTX_HD Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = mulhi32(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = mulhi32(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    Philox4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
    return o;
}

}  // namespace ptx
