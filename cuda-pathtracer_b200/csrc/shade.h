// shade.h -- per-path maths of the integrator, shared by the wavefront `shade` kernel, the fused
// bounce kernel and (compiled for the host) the CPU-side unit tests.
//
// Follows, line by line in behaviour, the BSDF-sampling branch of the reference integrator:
//   Sensor::get_ray + Ray ctor      rendering/sensor.h:31-33, core/ray.h:9-12
//   integrator()                    rendering/integrator.h:189-268
//   sampleCosineHemisphere()        rendering/integrator.h:62-85
//   tone map / quantise             rendering/integrator.h:393-407
// including the order in which random numbers are drawn: [depth > 2: 1 draw for
// Russian roulette] then 2 draws for the cosine sample, also at the last depth.
#pragma once
#include "layout.h"
#include "rng.h"

namespace ptx {

struct ShadeParams {
    int max_depth;        // reference: 5 (integrator.h:389)
    int rr_start_depth;   // reference: RR when depth > 2  => 3
    float rr_cap;         // reference: 0.95
};

struct PathVertex {       // what survives from one bounce to the next
    vec3 o, d;            // ray (d normalised)
    vec3 beta;            // throughput
    int depth;
};

PTX_HD void camera_ray(const CameraView& c, float u, float v, vec3& o, vec3& d) {
    o = ld3(c.origin);
    vec3 dir = ((ld3(c.llc) + u * ld3(c.horizontal)) + v * ld3(c.vertical)) - o;
    d = unit(dir);
}

PTX_HD vec3 cosine_hemisphere(vec3 n, float u, float v) {
    float r = sqrt_exact(u);
    float phi = (float)(2.0 * 3.14159265358979323846 * (double)v);   // `2.0f * M_PI * v` is double on Linux
    float sn, cs;
#if defined(__CUDA_ARCH__)
    sincosf(phi, &sn, &cs);
#else
    sn = sinf(phi); cs = cosf(phi);
#endif
    float x = r * cs, y = r * sn;
    float z = sqrt_exact(fmaxf(0.0f, 1.0f - u));
    vec3 tangent, bitangent;
    if (n.z < -0.9999999f) {
        tangent = mk3(0.0f, -1.0f, 0.0f);
        bitangent = mk3(-1.0f, 0.0f, 0.0f);
    } else {
        float a = 1.0f / (1.0f + n.z);
        float b = -n.x * n.y * a;
        tangent = mk3(1.0f - n.x * n.x * a, b, -n.x);
        bitangent = mk3(b, 1.0f - n.y * n.y * a, -n.y);
    }
    return unit((tangent * x + bitangent * y) + n * z);
}

// One bounce after `extend` found (prim, t).  Adds emitted radiance to `radiance`.
// Returns true if the path continues with the updated vertex, false if it ended.
// Control flow is deliberately single-exit (nested ifs, no early return): with early returns the
// compiler's reconvergence point moves to the end of the function and the lanes that took the
// Russian-roulette branch execute the rest of the shading separately from the others (ncu r1b:
// the shading block ran 1.8x per warp iteration at 38 % lane efficiency).
template <class Rng>
PTX_HD bool shade_bounce(const SceneView& s, const ShadeParams& sp, PathVertex& pv, uint32_t prim, float t, Rng& rng,
                         vec3& radiance) {
    bool cont = false;
    if (prim != kMiss) {                                                // miss: no environment light (integrator.h:199-201)
        f32x4 ps = ldq(s.prim_shade + prim);
        uint32_t mat = f2u(ps.w);
        f32x4 m0 = ldq(s.materials + 2 * (size_t)mat), m1 = ldq(s.materials + 2 * (size_t)mat + 1);
        vec3 Le = mk3(m1.x, m1.y, m1.z);
        radiance = radiance + pv.beta * Le;                            // both faces emit (integrator.h:204)
        bool survive = true;
        if (pv.depth >= sp.rr_start_depth) {                           // integrator.h:207-212
            float mx = fmaxf(pv.beta.x, fmaxf(pv.beta.y, pv.beta.z));
            float rr = fminf(mx, sp.rr_cap);
            survive = !(rng.next() > rr);
            if (survive) pv.beta = div_scalar(pv.beta, rr);
        }
        if (survive) {
            pv.beta = pv.beta * mk3(m0.x, m0.y, m0.z);
            if (!(length(pv.beta) < 1e-5f)) {                           // integrator.h:215-218
                vec3 n = mk3(ps.x, ps.y, ps.z);
                vec3 ns = dot(pv.d, n) < 0 ? n : -n;
                vec3 next;
                if (f2u(m0.w) == 1u) next = pv.d - (2.0f * dot(pv.d, ns)) * ns;     // mirror extension: no draws consumed
                else { float u = rng.next(); float v = rng.next(); next = cosine_hemisphere(ns, u, v); }
                vec3 p = pv.o + t * pv.d;                              // triangle.h:90
                pv.o = p + ns * 1e-4f;
                pv.d = unit(next);
                pv.depth++;
                cont = pv.depth < sp.max_depth;
            }
        }
    }
    return cont;
}

PTX_HD uint8_t tonemap_channel(float c) {
    c = c / (c + 1.0f);
    c = powf(c, 1.0f / 2.2f);
    return (uint8_t)(255.99f * fminf(c, 1.0f));
}

// RNG adaptors ---------------------------------------------------------------------------------
struct XorwowRng {
    Xorwow st;
    PTX_HD float next() { return xorwow_uniform(st); }
};
// Philox: one 4-wide block per (path, bounce); lanes handed out in order
struct PhiloxRng {
    Philox4 blk;
    int lane;
    PTX_HD void seek(uint64_t path, uint32_t block, uint64_t seed) {
        blk = philox4x32_10((uint32_t)path, (uint32_t)(path >> 32), block, 0u, (uint32_t)seed, (uint32_t)(seed >> 32));
        lane = 0;
    }
    PTX_HD float next() {
        uint32_t x = lane == 0 ? blk.x : (lane == 1 ? blk.y : (lane == 2 ? blk.z : blk.w));
        lane++;
        return u32_to_unit(x);
    }
};

}  // namespace ptx
