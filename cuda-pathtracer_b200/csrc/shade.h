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
    int sampling_mode;    // SamplingMode (render_config.h:41-47): 0 BSDF, 3 MIS, 1 / 2 / 4 pure grid sampling
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
    sincos_small(phi, sn, cs);          // == sincosf(phi, ...) for phi in (0, 2 pi], see common.h
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
        float a = rcp_exact(1.0f + n.z);
        float b = -n.x * n.y * a;
        tangent = mk3(1.0f - n.x * n.x * a, b, -n.x);
        bitangent = mk3(b, 1.0f - n.y * n.y * a, -n.y);
    }
    return unit((tangent * x + bitangent * y) + n * z);
}

// ---------------------------------------------------------------------------------------------------------------
// Grid-guided sampling: Grid (rendering/grid.h), PrecomputedCDF (render_config.h:27-34), sampleMIS (integrator.h:112-166).
// One guide record per reference primitive, kGuideFloats floats (the reference's PrecomputedCDF is 530: it also keeps the
// never-sampled lower hemisphere rows and the row sums):
//   [0,128)   pdf       luminance of the 8 x 16 upper-hemisphere cells (row = theta band, column = phi band)
//   [128,256) row_cdfs  conditional CDF of every row
//   [256,264) marginal  CDF over the rows
//   [264]     total_weight      [265] is_valid (0 / 1 as uint)      [266,272) padding (16-byte alignment)
// The reference evaluates every product with M_PI in DOUBLE (math.h's M_PI wins over grid.h:22-24's float fallback);
// kPiD reproduces that, rounding to float exactly where the C++ assigns to a float.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kGuideFloats = 272;
constexpr int kGuideRowCdfs = 128, kGuideMarginal = 256, kGuideTotal = 264, kGuideValid = 265;
constexpr double kPiD = 3.14159265358979323846;

// Guide record of one primitive from its 16 x 16 RGB radiosity grid: Grid::initFromRadiosity + buildCDFs (grid.h:62-137),
// the arithmetic SceneState::precomputeCDFs (application_state.h:492-585) runs once per radiosity solve.  Sums run in the
// reference's order; the luminance uses explicitly rounded products so that the device build equals the host build bit for bit.
PTX_HD float mul_rn(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fmul_rn(a, b);
#else
    return a * b;
#endif
}
PTX_HD float add_rn(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fadd_rn(a, b);
#else
    return a + b;
#endif
}
PTX_HD void build_guide_record(const float* rgb, float* out) {
    float total = 0.0f;
    for (int v = 0; v < 8; v++) {
        float row_sum = 0.0f;
        for (int u = 0; u < 16; u++) {
            const float* c = rgb + 3 * (v * 16 + u);
            float lum = add_rn(add_rn(mul_rn(0.2126f, c[0]), mul_rn(0.7152f, c[1])), mul_rn(0.0722f, c[2]));
            out[v * 16 + u] = lum;
            row_sum = add_rn(row_sum, lum);
        }
        out[kGuideRowCdfs + v * 16] = row_sum;          // parked here until the row's CDF is written below
        total = add_rn(total, row_sum);
    }
    const float inv_total = (total > 1e-6f) ? (1.0f / total) : 0.0f;
    float running = 0.0f;
    for (int v = 0; v < 8; v++) {
        const float row_sum = out[kGuideRowCdfs + v * 16];
        running = add_rn(running, row_sum);
        out[kGuideMarginal + v] = mul_rn(running, inv_total);
        if (row_sum < 1e-6f) {
            for (int u = 0; u < 16; u++) out[kGuideRowCdfs + v * 16 + u] = mul_rn((float)(u + 1), 0.0625f);
        } else {
            const float inv_row_sum = 1.0f / row_sum;
            float running_row = 0.0f;
            for (int u = 0; u < 16; u++) {
                running_row = add_rn(running_row, out[v * 16 + u]);
                out[kGuideRowCdfs + v * 16 + u] = mul_rn(running_row, inv_row_sum);
            }
            out[kGuideRowCdfs + v * 16 + 15] = 1.0f;
        }
    }
    out[kGuideMarginal + 7] = 1.0f;
    out[kGuideTotal] = total;
    out[kGuideValid] = u2f(total > 1e-6f ? 1u : 0u);
    for (int k = kGuideValid + 1; k < kGuideFloats; k++) out[k] = 0.0f;
}

PTX_HD float ldf(const float* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}
PTX_HD void sincos_exact(float a, float& sn, float& cs) {
#if defined(__CUDA_ARCH__)
    sincosf(a, &sn, &cs);
#else
    sn = sinf(a); cs = cosf(a);       // glibc's sincosf returns these very values
#endif
}
PTX_HD void frisvad_frame(vec3 n, vec3& t, vec3& b) {                       // grid.h:300-309 (= integrator.h:72-81)
    if (n.z < -0.9999999f) { t = mk3(0.0f, -1.0f, 0.0f); b = mk3(-1.0f, 0.0f, 0.0f); return; }
    const float a = 1.0f / (1.0f + n.z);
    const float c = -n.x * n.y * a;
    t = mk3(1.0f - n.x * n.x * a, c, -n.x);
    b = mk3(c, 1.0f - n.y * n.y * a, -n.y);
}
PTX_HD int guide_search(const float* cdf, int size, float xi) {             // linearSearchCDF, grid.h:248-255
    xi = fminf(fmaxf(xi, 0.0f), 0.999999f);
    int r = size - 1;
    for (int i = size - 1; i >= 0; i--) if (xi < ldf(cdf + i)) r = i;     // first i with xi < cdf[i]
    return r;
}
PTX_HD float guide_pdf_for_cell(const float* g, int theta_idx, int phi_idx) {   // computePDFForCell, grid.h:258-274
    float cell_value = ldf(g + theta_idx * 16 + phi_idx);
    if (cell_value < 1e-8f) return 1e-6f;
    float cell_prob = cell_value / fmaxf(ldf(g + kGuideTotal), 1e-6f);
    float theta_center = (float)((double)(((float)theta_idx + 0.5f) * 0.125f) * (kPiD * 0.5));
    float sin_theta = fmaxf(sinf(theta_center), 0.01f);
    float solid_angle = (float)(((double)sin_theta * ((kPiD * 0.5) / 8)) * (2.0 * kPiD / 16));
    return cell_prob / fmaxf(solid_angle, 1e-6f);
}
template <class Rng>
PTX_HD vec3 guide_sample(const float* g, vec3 normal, Rng& rng, float& out_pdf) {   // Grid::sample, grid.h:143-190 (valid grids)
    float xi1 = rng.next();
    float xi2 = rng.next();
    int theta_idx = guide_search(g + kGuideMarginal, 8, xi1);
    int phi_idx = guide_search(g + kGuideRowCdfs + theta_idx * 16, 16, xi2);
    float jitter_theta = rng.next();
    float jitter_phi = rng.next();
    float theta = (float)((double)(((float)theta_idx + jitter_theta) * 0.125f) * (kPiD * 0.5));
    theta = fminf(theta, (float)(kPiD * 0.5 - (double)0.01f));
    float phi = (float)((double)((((float)phi_idx + jitter_phi) * 0.0625f) * 2.0f) * kPiD);
    float sin_t, cos_t, sin_p, cos_p;
    sincos_exact(theta, sin_t, cos_t);
    sincos_exact(phi, sin_p, cos_p);
    vec3 tangent, bitangent;
    frisvad_frame(normal, tangent, bitangent);
    vec3 world = unit((tangent * (sin_t * cos_p) + bitangent * (sin_t * sin_p)) + normal * cos_t);
    out_pdf = guide_pdf_for_cell(g, theta_idx, phi_idx);
    return world;
}
PTX_HD float guide_compute_pdf(const float* g, vec3 dir, vec3 normal) {     // Grid::computePDF + worldToSpherical, grid.h:200-216, 312-325
    vec3 tangent, bitangent;
    frisvad_frame(normal, tangent, bitangent);
    float lx = dot(dir, tangent), ly = dot(dir, bitangent), lz = dot(dir, normal);
    float theta = acosf(fminf(fmaxf(lz, -1.0f), 1.0f));
    float phi = atan2f(ly, lx);
    if (phi < 0.0f) phi = (float)((double)phi + 2.0 * kPiD);
    if ((double)theta > kPiD * 0.5) return 0.0f;
    int theta_idx = (int)(((double)theta * (2.0 / kPiD)) * 8);
    int phi_idx = (int)(((double)phi * (0.5 / kPiD)) * 16);
    theta_idx = theta_idx < 0 ? 0 : (theta_idx > 7 ? 7 : theta_idx);
    phi_idx = phi_idx < 0 ? 0 : (phi_idx > 15 ? 15 : phi_idx);
    return guide_pdf_for_cell(g, theta_idx, phi_idx);
}
PTX_HD float mis_power_heuristic(float a, float b) {                         // integrator.h:90-95
    if (a <= 0.0f) return 0.0f;
    float a2 = a * a, b2 = b * b;
    return a2 / (a2 + b2);
}
// Next direction through the primitive's grid; `weight` multiplies the throughput (integrator.h:233-256).
// mode 3: one-sample MIS between the grid and the cosine lobe (sampleMIS); other modes: the grid alone.
template <class Rng>
PTX_HD vec3 guided_direction(const float* g, int mode, float mis_bsdf_fraction, vec3 ns, Rng& rng, float& weight) {
    vec3 dir;
    if (mode == 3) {
        const float BSDF_PROB = fmaxf(fminf(mis_bsdf_fraction, 0.99f), 0.01f);
        const float GRID_PROB = 1.0f - BSDF_PROB;
        float xi = rng.next();
        if (xi < BSDF_PROB) {
            float u = rng.next(); float v = rng.next();
            dir = cosine_hemisphere(ns, u, v);
            float cos_theta = fmaxf(dot(dir, ns), 0.0f);
            float pdf_bsdf = (float)((double)cos_theta / kPiD);
            float pdf_grid = guide_compute_pdf(g, dir, ns);
            float mis_w = mis_power_heuristic(pdf_bsdf, pdf_grid);
            weight = pdf_bsdf > 1e-6f ? mis_w / BSDF_PROB : 0.0f;
        } else {
            float pdf_grid;
            dir = guide_sample(g, ns, rng, pdf_grid);
            float cos_theta = fmaxf(dot(dir, ns), 0.0f);
            float pdf_bsdf = (float)((double)cos_theta / kPiD);
            float mis_w = mis_power_heuristic(pdf_grid, pdf_bsdf);
            if (pdf_grid > 1e-6f && cos_theta > 0.0f) {
                float w = (float)((double)(mis_w * cos_theta) / ((kPiD * (double)pdf_grid) * (double)GRID_PROB));
                weight = fminf(w, 10.0f);
            } else weight = 0.0f;
        }
    } else {
        float grid_pdf;
        dir = guide_sample(g, ns, rng, grid_pdf);
        float cos_theta = fmaxf(dot(dir, ns), 0.0f);
        float w = (float)((double)cos_theta / (kPiD * (double)fmaxf(grid_pdf, 1e-6f)));
        weight = fminf(fmaxf(w, 0.0f), 10.0f);
    }
    return dir;
}

// One bounce after `extend` found (prim, t).  Adds emitted radiance to `radiance`.
// Returns true if the path continues with the updated vertex, false if it ended.
// Control flow is deliberately single-exit (nested ifs, no early return): with early returns the
// compiler's reconvergence point moves to the end of the function and the lanes that took the
// Russian-roulette branch execute the rest of the shading separately from the others (ncu r1b:
// the shading block ran 1.8x per warp iteration at 38 % lane efficiency).
// GUIDED instantiations also honour sp.sampling_mode (grid / MIS modes when the scene carries guide records); the default
// instantiation contains no trace of that code.
// shade_hit: the hit is known to exist; n = the primitive's stored flat normal, mat = its material index.  shade_bounce (below)
// looks both up for an ordinary scene; the instanced kernel computes them per instance (traverse_inst.h resolve_instanced_hit).
template <class Rng, bool GUIDED = false>
PTX_HD bool shade_hit(const SceneView& s, const ShadeParams& sp, PathVertex& pv, uint32_t prim, vec3 n, uint32_t mat, float t, Rng& rng,
                      vec3& radiance) {
    bool cont = false;
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
            vec3 ns = dot(pv.d, n) < 0 ? n : -n;
            vec3 next;
            const float* g = nullptr;
            if (GUIDED && sp.sampling_mode != 0 && s.guide != nullptr) {        // initGridFromPrimitive + isValid, integrator.h:229-231
                g = s.guide + (size_t)prim * kGuideFloats;
                if (f2u(ldf(g + kGuideValid)) == 0u) g = nullptr;
            }
            if (f2u(m0.w) == 1u) next = pv.d - (2.0f * dot(pv.d, ns)) * ns;     // mirror extension: no draws consumed
            else if (GUIDED && g != nullptr) {
                float weight;
                next = guided_direction(g, sp.sampling_mode, s.mis_bsdf_fraction, ns, rng, weight);
                pv.beta = pv.beta * weight;
            } else { float u = rng.next(); float v = rng.next(); next = cosine_hemisphere(ns, u, v); }
            vec3 p = pv.o + t * pv.d;                              // triangle.h:90
            pv.o = p + ns * 1e-4f;
            pv.d = unit(next);
            pv.depth++;
            cont = pv.depth < sp.max_depth;
        }
    }
    return cont;
}

template <class Rng, bool GUIDED = false>
PTX_HD bool shade_bounce(const SceneView& s, const ShadeParams& sp, PathVertex& pv, uint32_t prim, float t, Rng& rng,
                         vec3& radiance) {
    bool cont = false;
    if (prim != kMiss) {                                                // miss: no environment light (integrator.h:199-201)
        f32x4 ps = ldq(s.prim_shade + prim);
        cont = shade_hit<Rng, GUIDED>(s, sp, pv, prim, mk3(ps.x, ps.y, ps.z), f2u(ps.w), t, rng, radiance);
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
