// kernels.cu -- hand-written sm_100a kernels of the wavefront integrator (see kernels.cuh).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3
#include "kernels.cuh"

#include <mutex>
#include <type_traits>
#include "traverse.h"
#include "traverse_inst.h"
#include "flat.h"

namespace ptx {

namespace {

constexpr int kBlock = 256;

__device__ __forceinline__ float4 ld_state(const float4* p) { return *p; }
__device__ __forceinline__ vec3 v3(float x, float y, float z) { return mk3(x, y, z); }

// one atomicAdd per warp for a per-lane 0/1 contribution
__device__ __forceinline__ void warp_count_add(unsigned long long* ctr, bool pred) {
    unsigned m = __ballot_sync(0xFFFFFFFFu, pred);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(ctr, (unsigned long long)__popc(m));
}

// Pixel hand-off between the sample chunks of one pixel (compat RNG, k_path_persist / k_path_flat): the lane that finished
// chunk c stores the pixel's RNG state and running sum, then RELEASES progress[pixel] = c + 1; the lane holding chunk c + 1
// polls the flag and reads the state.
//   st_progress    : st.release.gpu -- orders the state stores before the flag.
//   progress_ready : polls with ld.relaxed.gpu (served by L2, no side effect) and, ONCE the expected value is seen, reads the
//                    flag again with ld.acquire.gpu.  That acquire load observes the release store (or a later value in the
//                    flag's modification order -- there is none before this lane itself writes c + 2), so the state loads
//                    that follow it are ordered after the previous owner's stores by the PTX memory model itself
//                    (release / acquire on one location), not by a control dependency plus the observation that .cg loads
//                    bypass L1, as in round 1.  ptxas compiles ld.acquire.gpu to LDG.E.STRONG.GPU + CCTL.IVALL, i.e. it
//                    invalidates the SM's L1 (seen in the SASS and as a stall site in profiles/r2b_*): paying that on every
//                    poll cost 6 % on the BVH8 kernel, whose tree lives in L1; once per chunk start it is 6 invalidations
//                    per pixel and frame.  -DPTX_HANDOFF_RELAXED builds the round-1 protocol for A/B measurements.
__device__ __forceinline__ bool progress_ready(const uint32_t* p, uint32_t expect) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    if (v != expect) return false;
#if !defined(PTX_HANDOFF_RELAXED)
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
#endif
    return v == expect;
}
__device__ __forceinline__ void st_progress(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// ---------------------------------------------------------------------------------------------
// rng_init: XORWOW state of pixel p == curand_init(seed_base + p, p, 0)   (integrator.h:274-280)
// The jump J[b] is applied for every set bit b of p; the matrix rows are read with warp-uniform
// addresses (broadcast loads), the per-lane state bits select them with masks.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_rng_init(FilmView film, unsigned long long seed_base, const uint32_t* __restrict__ jump) {
    uint32_t p = blockIdx.x * kBlock + threadIdx.x;
    uint32_t npix = (uint32_t)film.width * (uint32_t)film.height;
    bool valid = p < npix;
    Xorwow s = xorwow_seed(seed_base + (unsigned long long)p);
    uint32_t top = __reduce_max_sync(0xFFFFFFFFu, valid ? p : 0u);   // highest pixel index in the warp
    for (int b = 0; b < 32 && (top >> b) != 0; b++) {
        if (__any_sync(0xFFFFFFFFu, valid && ((p >> b) & 1u))) {
            Xorwow t = s;
            xorwow_apply(t, jump + (size_t)b * kXorwowJumpWords);
            if ((p >> b) & 1u) s = t;
        }
    }
    if (valid) {
        film.rngA[p] = make_uint4(s.v0, s.v1, s.v2, s.v3);
        film.rngB[p] = make_uint2(s.v4, s.d);
    }
}

// slot -> pixel map of the pixels this rank owns (tile sharding deals bands of kBandRows rows)
__global__ void __launch_bounds__(kBlock) k_slot_pixels(uint32_t* slot_pixel, int w, int h, int shard_mode, int rank, int world,
                                                        uint32_t n_owned) {
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= n_owned) return;
    if (shard_mode != 1 || world <= 1) { slot_pixel[i] = i; return; }
    uint32_t band_px = (uint32_t)kBandRows * (uint32_t)w;
    uint32_t local_band = i / band_px, within = i % band_px;
    uint32_t band = local_band * (uint32_t)world + (uint32_t)rank;
    slot_pixel[i] = band * band_px + within;
}

// Whole-pixel culling: see launch_cull_pixels (kernels.cuh).  A culled pixel's samples all miss the scene (its rays lie outside the
// projection of the scene's bounds: api.cu cull_rect), so each consumes exactly the two jitter draws (integrator.h:384-385) and adds
// nothing to the film: the pixel's stream is advanced by 2 * spp draws with one GF(2) matrix product and its running sum is left alone
// -- the film and the RNG state are what tracing the pixel leaves behind, bit for bit.
__global__ void __launch_bounds__(kBlock) k_cull_pixels(FilmView film, const uint32_t* __restrict__ slot_base, uint32_t n_owned, CullRect r,
                                                        const uint32_t* __restrict__ row_prefix, const uint32_t* __restrict__ skip, uint32_t weyl_add,
                                                        uint32_t* __restrict__ slot_active) {
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= n_owned) return;
    const uint32_t p = slot_base[i];
    const int x = (int)(p % (uint32_t)film.width), y = (int)(p / (uint32_t)film.width);
    if (x >= r.x0 && x < r.x1 && y >= r.y0 && y < r.y1) {
        slot_active[row_prefix[y] * (uint32_t)(r.x1 - r.x0) + (uint32_t)(x - r.x0)] = p;
        return;
    }
    if (skip == nullptr) return;
    uint4 a = film.rngA[p]; uint2 b = film.rngB[p];
    Xorwow s{a.x, a.y, a.z, a.w, b.x, b.y};
    xorwow_apply(s, skip);
    s.d += weyl_add;
    film.rngA[p] = make_uint4(s.v0, s.v1, s.v2, s.v3);
    film.rngB[p] = make_uint2(s.v4, s.d);
}

__device__ __forceinline__ void store_vertex(const WaveState& ws, uint32_t slot, const PathVertex& pv, uint32_t flags) {
    ws.A[slot] = make_float4(pv.o.x, pv.o.y, pv.o.z, pv.d.x);
    ws.B[slot] = make_float4(pv.d.y, pv.d.z, __uint_as_float((uint32_t)pv.depth | flags), pv.beta.x);
    float4 c = ws.C[slot];
    c.x = pv.beta.y; c.y = pv.beta.z;
    ws.C[slot] = c;
}

__device__ __forceinline__ void start_path(const CameraView& cam, uint32_t pixel, float ju, float jv, PathVertex& pv) {
    int x = (int)(pixel % (uint32_t)cam.width), y = (int)(pixel / (uint32_t)cam.width);
    float u = ((float)x + ju) / (float)cam.width;            // integrator.h:384-385
    float v = ((float)y + jv) / (float)cam.height;
    camera_ray(cam, u, v, pv.o, pv.d);
    pv.beta = mk3(1.0f, 1.0f, 1.0f);
    pv.depth = 0;
}

// ---------------------------------------------------------------------------------------------
// raygen: first path of every slot
// ---------------------------------------------------------------------------------------------
template <int RNG>
__global__ void __launch_bounds__(kBlock) k_raygen(WaveState ws, FilmView film, CameraView cam, RenderParams rp) {
    uint32_t slot = blockIdx.x * kBlock + threadIdx.x;
    if (slot >= rp.n_slots) return;
    ws.queue[0][slot] = slot;
    PathVertex pv;
    uint32_t flags = 0;
    if (RNG == 0) {
        uint32_t pixel = ws.slot_pixel[slot];
        uint4 a = film.rngA[pixel]; uint2 b = film.rngB[pixel];
        XorwowRng rng; rng.st = Xorwow{a.x, a.y, a.z, a.w, b.x, b.y};
        if (rp.spp > 0) {
            float ju = rng.next(), jv = rng.next();
            start_path(cam, pixel, ju, jv, pv);
            film.rngA[pixel] = make_uint4(rng.st.v0, rng.st.v1, rng.st.v2, rng.st.v3);
            film.rngB[pixel] = make_uint2(rng.st.v4, rng.st.d);
        } else { pv.o = pv.d = pv.beta = mk3(0, 0, 0); pv.depth = 0; flags = kDeadFlag; }
        ws.D[slot] = make_uint2(pixel, 0u);
    } else {
        // path index == slot for the first wave; pixel-major so neighbouring lanes take neighbouring pixels
        unsigned long long path = slot;
        if (path < rp.total_paths) {
            uint32_t local = (uint32_t)(path % rp.n_owned), sample = (uint32_t)(path / rp.n_owned) + rp.sample_begin;
            uint32_t pixel = ws.slot_pixel[local];
            PhiloxRng rng; rng.seek(((unsigned long long)sample << 32) | pixel, 0u, rp.seed);
            float ju = rng.next(), jv = rng.next();
            start_path(cam, pixel, ju, jv, pv);
            ws.D[slot] = make_uint2(pixel, sample);
        } else { pv.o = pv.d = pv.beta = mk3(0, 0, 0); pv.depth = 0; flags = kDeadFlag; ws.D[slot] = make_uint2(0u, 0u); }
    }
    ws.C[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
    ws.E[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
    store_vertex(ws, slot, pv, flags);
}


// ---------------------------------------------------------------------------------------------
// extend: closest hit for the ray of every live slot in the queue
// ---------------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(kBlock) k_extend(WaveState ws, SceneView scene, int q) {
    uint32_t k = blockIdx.x * kBlock + threadIdx.x;
    uint32_t qlen = (uint32_t)ws.counters[CTR_QLEN];
    if (k >= qlen) return;
    uint32_t slot = ws.queue[q][k];
    float4 a = ws.A[slot], b = ws.B[slot];
    if (__float_as_uint(b.z) & kDeadFlag) return;
    float t;
    TraceCounters tc = {0u, 0u};
    uint32_t tri = trace_closest<24, COUNT>(scene, v3(a.x, a.y, a.z), v3(a.w, b.x, b.y), 1e-4f, FLT_MAX, t, &tc);
    uint32_t prim = kMiss;
    if (tri != kMiss) prim = __float_as_uint(ldq(scene.tris + (size_t)tri * kTriQuads).w);
    reinterpret_cast<float2*>(ws.C + slot)[1] = make_float2(t, __uint_as_float(prim));
    if (COUNT) { atomicAdd(&ws.counters[CTR_NODE_VISITS], (unsigned long long)tc.nodes); atomicAdd(&ws.counters[CTR_TRI_TESTS], (unsigned long long)tc.tris); }
}

// ---------------------------------------------------------------------------------------------
// shade (+ Russian roulette + film accumulate + in-place regeneration).  FUSED = true also does the
// extend step in the same kernel ("bounce" pipeline), so no hit record goes through HBM.
// ---------------------------------------------------------------------------------------------
template <int RNG, bool FUSED>
__global__ void __launch_bounds__(kBlock) k_shade(WaveState ws, SceneView scene, FilmView film, CameraView cam, RenderParams rp, int q) {
    uint32_t k = blockIdx.x * kBlock + threadIdx.x;
    uint32_t qlen = (uint32_t)ws.counters[CTR_QLEN];
    bool live = k < qlen;
    uint32_t slot = live ? ws.queue[q][k] : 0u;
    float4 a, b, c;
    if (live) {
        b = ws.B[slot];
        live = !(__float_as_uint(b.z) & kDeadFlag);
    }
    bool ended = false, died = false;
    if (live) {
        a = ws.A[slot]; c = ws.C[slot];
        PathVertex pv;
        pv.o = v3(a.x, a.y, a.z); pv.d = v3(a.w, b.x, b.y);
        pv.beta = v3(b.w, c.x, c.y);
        pv.depth = (int)(__float_as_uint(b.z) & 0xFFFFu);
        uint2 dd = ws.D[slot];
        uint32_t pixel = dd.x;
        float t = c.z; uint32_t prim = __float_as_uint(c.w);
        if (FUSED) {
            uint32_t tri = trace_closest<24, false>(scene, pv.o, pv.d, 1e-4f, FLT_MAX, t, nullptr);
            prim = tri == kMiss ? kMiss : __float_as_uint(ldq(scene.tris + (size_t)tri * kTriQuads).w);
        }
        vec3 rad = mk3(0.f, 0.f, 0.f);
        bool cont;
        XorwowRng xr; PhiloxRng pr;
        if (RNG == 0) {
            uint4 ra = film.rngA[pixel]; uint2 rb = film.rngB[pixel];
            xr.st = Xorwow{ra.x, ra.y, ra.z, ra.w, rb.x, rb.y};
            cont = shade_bounce(scene, rp.shade, pv, prim, t, xr, rad);
        } else {
            pr.seek(((unsigned long long)dd.y << 32) | pixel, 1u + (uint32_t)pv.depth, rp.seed);
            cont = shade_bounce(scene, rp.shade, pv, prim, t, pr, rad);
        }
        if (rad.x != 0.f || rad.y != 0.f || rad.z != 0.f) {
            if (RNG == 0) {            // the pixel is private to this slot: plain read-modify-write
                float4 acc = film.accum[pixel];
                acc.x += rad.x; acc.y += rad.y; acc.z += rad.z;
                film.accum[pixel] = acc;
            } else {                   // samples of one pixel may be in flight in several slots
                float* acc = reinterpret_cast<float*>(film.accum + pixel);
                atomicAdd(acc + 0, rad.x); atomicAdd(acc + 1, rad.y); atomicAdd(acc + 2, rad.z);
            }
        }
        ended = !cont;
        uint32_t flags = 0;
        if (RNG == 0) {
            if (ended) {
                uint32_t done = dd.y + 1u;
                if (done < (uint32_t)rp.spp) {
                    float ju = xr.next(), jv = xr.next();
                    start_path(cam, pixel, ju, jv, pv);
                } else { died = true; flags = kDeadFlag; }
                ws.D[slot] = make_uint2(pixel, done);
            }
            film.rngA[pixel] = make_uint4(xr.st.v0, xr.st.v1, xr.st.v2, xr.st.v3);
            film.rngB[pixel] = make_uint2(xr.st.v4, xr.st.d);
        }
        if (RNG == 1) {
            // warp-aggregated fetch of the next path indices (one atomic per warp)
            unsigned need = __ballot_sync(__activemask(), ended);
            if (ended) {
                int leader = __ffs(need) - 1;
                unsigned long long base = 0;
                if ((int)(threadIdx.x & 31) == leader) base = atomicAdd(&ws.counters[CTR_NEXT_PATH], (unsigned long long)__popc(need));
                base = __shfl_sync(need, base, leader);
                unsigned long long path = base + (unsigned long long)__popc(need & ((1u << (threadIdx.x & 31)) - 1u));
                if (path < rp.total_paths) {
                    uint32_t local = (uint32_t)(path % rp.n_owned), sample = (uint32_t)(path / rp.n_owned) + rp.sample_begin;
                    uint32_t npx = ws.slot_pixel[local];
                    pr.seek(((unsigned long long)sample << 32) | npx, 0u, rp.seed);
                    float ju = pr.next(), jv = pr.next();
                    start_path(cam, npx, ju, jv, pv);
                    ws.D[slot] = make_uint2(npx, sample);
                } else { died = true; flags = kDeadFlag; }
            }
        }
        ws.A[slot] = make_float4(pv.o.x, pv.o.y, pv.o.z, pv.d.x);
        ws.B[slot] = make_float4(pv.d.y, pv.d.z, __uint_as_float((uint32_t)pv.depth | flags), pv.beta.x);
        if (FUSED) ws.C[slot] = make_float4(pv.beta.y, pv.beta.z, 0.f, 0.f);
        else reinterpret_cast<float2*>(ws.C + slot)[0] = make_float2(pv.beta.y, pv.beta.z);
    }
    // statistics: one atomic per warp per counter
    unsigned m_live = __ballot_sync(0xFFFFFFFFu, live);
    unsigned m_end = __ballot_sync(0xFFFFFFFFu, ended);
    unsigned m_died = __ballot_sync(0xFFFFFFFFu, died);
    if ((threadIdx.x & 31) == 0) {
        if (m_live) atomicAdd(&ws.counters[CTR_SEGMENTS], (unsigned long long)__popc(m_live));
        if (m_end) atomicAdd(&ws.counters[CTR_PATHS], (unsigned long long)__popc(m_end));
        if (m_died) atomicAdd(&ws.counters[CTR_LIVE], (unsigned long long)(-(long long)__popc(m_died)));
    }
}


// ---------------------------------------------------------------------------------------------
// path loop: the default pipeline.  One thread owns one slot for up to `iters` bounces: extend (BVH8
// closest hit), shade (emission, Russian roulette, albedo, cosine sample) and in-place regeneration of
// ended paths all run with the path state, the RNG state and the pixel's running sum in registers.
// The SoA state arrays are read once on entry and written once on exit, so per-bounce HBM traffic
// drops from ~150 B to ~(150 + 48) / iters B.  Slots whose pixel ran out of samples leave the loop;
// the host compacts the queue between launches (k_compact_*).
// Summation order equals the reference's: L accumulates along a path (integrator.h:204) and is added
// to the pixel sum when the path ends (`color += integrator(...)`, integrator.h:387-390).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void start_path_xy(const CameraView& cam, float fx, float fy, float ju, float jv, PathVertex& pv) {
    float u = div_exact_pos(fx + ju, (float)cam.width);      // integrator.h:384-385
    float v = div_exact_pos(fy + jv, (float)cam.height);
    camera_ray(cam, u, v, pv.o, pv.d);
    pv.beta = mk3(1.0f, 1.0f, 1.0f);
    pv.depth = 0;
}

template <int RNG, int MINB>
__global__ void __launch_bounds__(kBlock, MINB) k_path_loop(WaveState ws, SceneView scene, FilmView film, CameraView cam, RenderParams rp, int q,
                                                      int iters) {
    constexpr unsigned kFull = 0xFFFFFFFFu;
    uint32_t k = blockIdx.x * kBlock + threadIdx.x;
    uint32_t qlen = (uint32_t)ws.counters[CTR_QLEN];
    bool live = k < qlen;
    uint32_t slot = live ? ws.queue[q][k] : 0u;
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live) {
        b = ws.B[slot];
        live = !(__float_as_uint(b.z) & kDeadFlag);
    }
    uint32_t nseg = 0, npaths = 0, ndied = 0;
    const uint32_t nlive = __popc(__ballot_sync(kFull, live));
    PathVertex pv;
    pv.o = pv.d = pv.beta = v3(0.f, 0.f, 0.f); pv.depth = 0;
    vec3 L = v3(0.f, 0.f, 0.f), acc = v3(0.f, 0.f, 0.f);
    uint32_t pixel = 0, sample = 0;                     // compat: sample = samples finished; philox: global sample index
    XorwowRng xr; xr.st = Xorwow{0u, 0u, 0u, 0u, 0u, 0u};
    PhiloxRng pr; pr.blk = Philox4{0u, 0u, 0u, 0u}; pr.lane = 0;
    float fx = 0.f, fy = 0.f;
    if (live) {
        float4 a = ws.A[slot], c = ws.C[slot], e = ws.E[slot];
        uint2 dd = ws.D[slot];
        pv.o = v3(a.x, a.y, a.z); pv.d = v3(a.w, b.x, b.y);
        pv.beta = v3(b.w, c.x, c.y);
        pv.depth = (int)(__float_as_uint(b.z) & 0xFFFFu);
        L = v3(e.x, e.y, e.z);
        pixel = dd.x; sample = dd.y;
        if (RNG == 0) {
            uint4 ra = film.rngA[pixel]; uint2 rb = film.rngB[pixel];
            xr.st = Xorwow{ra.x, ra.y, ra.z, ra.w, rb.x, rb.y};
            float4 f = film.accum[pixel];
            acc = v3(f.x, f.y, f.z);
            fx = (float)(int)(pixel % (uint32_t)cam.width); fy = (float)(int)(pixel / (uint32_t)cam.width);
        }
    }
    // Per-lane state machine.  Every iteration each active lane advances its ray by ONE wide-node visit
    // (Traversal::step: node test + the triangles of the leaves it hit); lanes whose traversal completed
    // shade, regenerate if the path ended, and start the next ray at once, instead of idling until the
    // slowest ray of the warp is through (ncu r1b: 2.9 node blocks / 12.7 triangle blocks per warp and
    // segment at 54 % / 30 % lane efficiency with the ray-at-a-time loop).  The whole warp stays in the
    // loop so that __syncwarp can force reconvergence between the phases.  After `iters` steps lanes
    // park at their next segment boundary (the traversal stack is not saved across launches).
    bool alive = live, parked = false;
    Traversal tr;
    TravStack<24> stk;
    tr.begin(pv.d, FLT_MAX);
    for (int it = 0;; it++) {
        const bool active = alive && !parked;
        if (!__any_sync(kFull, active)) break;
        bool done = false;
        if (active) done = tr.template step<false, 24>(scene, stk, pv.o, pv.d, 1e-4f, nullptr);
        __syncwarp();
        bool cont = true;
        if (done) {
            uint32_t prim = tr.best == kMiss ? kMiss : __float_as_uint(ldq(scene.tris + (size_t)tr.best * kTriQuads).w);
            nseg++;
            if (RNG == 0) cont = shade_bounce(scene, rp.shade, pv, prim, tr.closest, xr, L);
            else {
                pr.seek(((unsigned long long)sample << 32) | pixel, 1u + (uint32_t)pv.depth, rp.seed);
                cont = shade_bounce(scene, rp.shade, pv, prim, tr.closest, pr, L);
            }
        }
        __syncwarp();
        if (done && !cont) {                             // path ended: commit it, start the pixel's next sample or die
            npaths++;
            if (RNG == 0) {
                acc = acc + L;
                sample++;
                if (sample < (uint32_t)rp.spp) {
                    float ju = xr.next(), jv = xr.next();
                    start_path_xy(cam, fx, fy, ju, jv, pv);
                } else alive = false;
            } else {
                if (L.x != 0.f || L.y != 0.f || L.z != 0.f) {      // samples of one pixel are in flight in several slots
                    float* fa = reinterpret_cast<float*>(film.accum + pixel);
                    atomicAdd(fa + 0, L.x); atomicAdd(fa + 1, L.y); atomicAdd(fa + 2, L.z);
                }
                // warp-aggregated fetch of the next path index: one atomic for all lanes that ended this iteration
                unsigned need = __activemask();
                int leader = __ffs(need) - 1, lane = (int)(threadIdx.x & 31);
                unsigned long long base = 0;
                if (lane == leader) base = atomicAdd(&ws.counters[CTR_NEXT_PATH], (unsigned long long)__popc(need));
                base = __shfl_sync(need, base, leader);
                unsigned long long path = base + (unsigned long long)__popc(need & ((1u << lane) - 1u));
                if (path < rp.total_paths) {
                    uint32_t local = (uint32_t)(path % rp.n_owned);
                    sample = (uint32_t)(path / rp.n_owned) + rp.sample_begin;
                    pixel = ws.slot_pixel[local];
                    pr.seek(((unsigned long long)sample << 32) | pixel, 0u, rp.seed);
                    float ju = pr.next(), jv = pr.next();
                    start_path_xy(cam, (float)(int)(pixel % (uint32_t)cam.width), (float)(int)(pixel / (uint32_t)cam.width), ju, jv, pv);
                } else alive = false;
            }
            L = v3(0.f, 0.f, 0.f);
        }
        __syncwarp();
        if (done) {
            if (it >= iters) parked = true;
            else tr.begin(pv.d, FLT_MAX);
        }
    }
    if (live) {
        uint32_t flags = alive ? 0u : kDeadFlag;
        ndied = alive ? 0u : 1u;
        ws.A[slot] = make_float4(pv.o.x, pv.o.y, pv.o.z, pv.d.x);
        ws.B[slot] = make_float4(pv.d.y, pv.d.z, __uint_as_float((uint32_t)pv.depth | flags), pv.beta.x);
        ws.C[slot] = make_float4(pv.beta.y, pv.beta.z, 0.f, 0.f);
        ws.E[slot] = make_float4(L.x, L.y, L.z, 0.f);
        ws.D[slot] = make_uint2(pixel, sample);
        if (RNG == 0) {
            film.rngA[pixel] = make_uint4(xr.st.v0, xr.st.v1, xr.st.v2, xr.st.v3);
            film.rngB[pixel] = make_uint2(xr.st.v4, xr.st.d);
            film.accum[pixel] = make_float4(acc.x, acc.y, acc.z, 0.f);
        }
    }
    // statistics: one atomic per warp per counter
    nseg = __reduce_add_sync(0xFFFFFFFFu, nseg);
    npaths = __reduce_add_sync(0xFFFFFFFFu, npaths);
    ndied = __reduce_add_sync(0xFFFFFFFFu, ndied);
    if ((threadIdx.x & 31) == 0) {
        if (nseg) atomicAdd(&ws.counters[CTR_SEGMENTS], (unsigned long long)nseg);
        if (npaths) atomicAdd(&ws.counters[CTR_PATHS], (unsigned long long)npaths);
        if (ndied) atomicAdd(&ws.counters[CTR_LIVE], (unsigned long long)(-(long long)ndied));
        if (nlive) atomicAdd(&ws.counters[CTR_SLOT_LAUNCHES], (unsigned long long)nlive);
    }
}

// guide records (shade.h) from per-primitive 16 x 16 RGB radiosity grids: SceneState::precomputeCDFs on the device
__global__ void __launch_bounds__(kBlock) k_build_guide(const float* __restrict__ rgb, size_t n_prims, float* __restrict__ out) {
    size_t p = (size_t)blockIdx.x * kBlock + threadIdx.x;
    if (p < n_prims) build_guide_record(rgb + p * 768, out + p * kGuideFloats);
}

// ---------------------------------------------------------------------------------------------
// persistent path loop (pipeline 3): ONE launch per frame, one resident CTA set (SMs x MINB), no path
// state in memory at all.  A lane pulls a work item from a global counter, carries it to the end in
// registers and pulls the next one:
//   compat RNG : item = one owned pixel.  The lane loads the pixel's XORWOW state and running sum, traces
//                its spp samples one after the other (the reference's per-pixel stream order,
//                integrator.h:370-392) and writes both back.
//   counter RNG: item = one path (sample s of pixel p); radiance is added to the film with atomics.
// Compared with k_path_loop this removes the per-launch state round trip, the queue compaction and --
// the big one -- the lanes that idle from the moment their pixel runs out of samples until the launch
// ends (ncu r1h: lane efficiency 0.53 in the second launch of a 128-spp frame).  Lanes only idle in the
// tail of the frame, when the counter has run dry.
// ---------------------------------------------------------------------------------------------
// INST: the scene is two-level (traverse_inst.h): the cursor is an InstTraversal, normal and material of a hit are resolved per instance.
template <int RNG, int BLOCK, int MINB, bool GUIDED = false, bool PAIR = false, bool INST = false>
__global__ void __launch_bounds__(BLOCK, MINB) k_path_persist(WaveState ws, SceneView scene, FilmView film, CameraView cam, RenderParams rp) {
    constexpr unsigned kFull = 0xFFFFFFFFu;
    const int lane = (int)(threadIdx.x & 31);
    // compat: item = (chunk c, owned pixel k), chunk-major, so chunk c of a pixel is handed out long after chunk c-1.
    // Chunk c covers samples [chunk_start[c], chunk_start[c+1]) -- the host picks the boundaries (api.cu: equal parts whose last
    // one is halved down to spp / 32), so most samples pay no hand-off and the frame's tail is bounded by the small last chunks.
    const uint32_t n_chunks = RNG == 0 ? (uint32_t)rp.n_chunks : 0u;
    const unsigned long long n_items = RNG == 0 ? (unsigned long long)rp.n_owned * n_chunks : rp.total_paths;
    uint32_t nseg = 0, npaths = 0;
    PathVertex pv;
    pv.o = pv.d = pv.beta = v3(0.f, 0.f, 0.f); pv.depth = 0;
    vec3 L = v3(0.f, 0.f, 0.f), acc = v3(0.f, 0.f, 0.f);
    uint32_t pixel = 0, sample = 0;                     // compat: samples finished; philox: global sample index
    uint32_t chunk = 0;                                 // compat: chunk number of the item in flight
    XorwowRng xr; xr.st = Xorwow{0u, 0u, 0u, 0u, 0u, 0u};
    PhiloxRng pr; pr.blk = Philox4{0u, 0u, 0u, 0u}; pr.lane = 0;
    float fx = 0.f, fy = 0.f;
    bool alive = false, pending = false, dry = rp.spp <= 0;
    bool waiting = false;                               // the lane's ray is through, its result not shaded yet
    typename std::conditional<INST, InstTraversal, Traversal>::type tr;
    TravStack<24> stk;
    auto restart = [&]() { if constexpr (INST) tr.begin(pv.o, pv.d, FLT_MAX); else tr.begin(pv.d, FLT_MAX); };
    restart();
    for (;;) {
        // ---- fetch: every lane without work asks for an item; one atomic per warp
        const bool want = !alive && !pending && !dry;
        const unsigned need = __ballot_sync(kFull, want);
        if (need) {
            unsigned long long base = 0;
            const int leader = __ffs(need) - 1;
            if (lane == leader) base = atomicAdd(&ws.counters[CTR_NEXT_PATH], (unsigned long long)__popc(need));
            base = __shfl_sync(kFull, base, leader);
            if (want) {
                const unsigned long long item = base + (unsigned long long)__popc(need & ((1u << lane) - 1u));
                if (item < n_items) {
                    if (RNG == 0) {
                        chunk = (uint32_t)(item / rp.n_owned);
                        pixel = ws.slot_pixel[(uint32_t)(item % rp.n_owned)];
                        pending = true;
                    } else {
                        pixel = ws.slot_pixel[(uint32_t)(item % rp.n_owned)];
                        sample = (uint32_t)(item / rp.n_owned) + rp.sample_begin;
                        pr.seek(((unsigned long long)sample << 32) | pixel, 0u, rp.seed);
                        float ju = pr.next(), jv = pr.next();
                        start_path_xy(cam, (float)(int)(pixel % (uint32_t)cam.width), (float)(int)(pixel / (uint32_t)cam.width), ju, jv, pv);
                        L = v3(0.f, 0.f, 0.f);
                        restart();
                        alive = true;
                    }
                } else dry = true;
            }
        }
        // ---- compat: start the chunk once its predecessor has been committed (almost always at once; a lane that has to
        //      wait just retries next iteration, so lanes of one warp can never block each other)
        if (RNG == 0 && pending) {
            bool ready = chunk == 0u;
            if (!ready) {
                ready = progress_ready(ws.progress + pixel, chunk);            // relaxed poll, one acquire load on success
            }
            if (ready) {
                // The flag was read with ld.acquire (progress_ready): the loads below are ordered after it and see what the previous
                // owner stored before its release.  They go to L2 (.cg) because an L1 line filled earlier could be stale; no
                // gpu-scope FENCE on purpose: that invalidates the SM's whole L1 (CCTL.IVALL), which holds the BVH (-20 % on C2).
                uint4 ra; uint2 rb; float4 f;
                if (chunk == 0u) { ra = film.rngA[pixel]; rb = film.rngB[pixel]; f = film.accum[pixel]; }
                else { ra = __ldcg(film.rngA + pixel); rb = __ldcg(film.rngB + pixel); f = __ldcg(film.accum + pixel); }   // L2: L1 may hold a stale line
                xr.st = Xorwow{ra.x, ra.y, ra.z, ra.w, rb.x, rb.y};
                acc = v3(f.x, f.y, f.z);
                sample = rp.chunk_start[chunk];
                fx = (float)(int)(pixel % (uint32_t)cam.width); fy = (float)(int)(pixel / (uint32_t)cam.width);
                float ju = xr.next(), jv = xr.next();
                start_path_xy(cam, fx, fy, ju, jv, pv);
                L = v3(0.f, 0.f, 0.f);
                restart();
                pending = false; alive = true;
            }
        }
        if (!__any_sync(kFull, alive || pending)) break;
        // ---- extend: one wide-node visit (+ its triangles) per lane
        if (alive && !waiting) {
            if constexpr (INST) waiting = tr.template step<24>(scene, stk, pv.o, pv.d, 1e-4f);
            else waiting = tr.template step<false, 24, PAIR>(scene, stk, pv.o, pv.d, 1e-4f, nullptr);
        }
        __syncwarp();
        // ---- shade in batches: a finished ray waits until enough lanes of the warp are through.  Shading + regeneration
        //      cost about as much as a node visit and run for the whole warp whenever one lane needs them; on a scene with
        //      n node visits per ray only 32 / n lanes finish per iteration (C2 21, C3 8, a C5-shaped scene 1), so shading at
        //      once wastes up to 31 / 32 of that block.  Waiting costs the idle lane's share of the node visits instead; the
        //      minimum of the two is near `shade_batch` = 12 waiting lanes for all three scenes (fewer when few lanes are alive).
        const int n_wait = __popc(__ballot_sync(kFull, waiting)), n_alive = __popc(__ballot_sync(kFull, alive));
        if (n_wait == 0 || n_wait < min(rp.shade_batch, (3 * n_alive + 7) >> 3)) continue;
        const bool done = waiting;
        waiting = false;
        bool cont = true;
        if (done) {
            nseg++;
            if constexpr (INST) {
                cont = false;
                if (tr.tr.best != kMiss) {
                    uint32_t prim, mat; vec3 n;
                    resolve_instanced_hit(scene, tr, prim, n, mat);
                    if (RNG == 0) cont = shade_hit<XorwowRng, false>(scene, rp.shade, pv, prim, n, mat, tr.tr.closest, xr, L);
                    else {
                        pr.seek(((unsigned long long)sample << 32) | pixel, 1u + (uint32_t)pv.depth, rp.seed);
                        cont = shade_hit<PhiloxRng, false>(scene, rp.shade, pv, prim, n, mat, tr.tr.closest, pr, L);
                    }
                }
            } else {
                uint32_t prim = tr.bestw;      // kMiss when nothing was hit
                if (RNG == 0) cont = shade_bounce<XorwowRng, GUIDED>(scene, rp.shade, pv, prim, tr.closest, xr, L);
                else {
                    pr.seek(((unsigned long long)sample << 32) | pixel, 1u + (uint32_t)pv.depth, rp.seed);
                    cont = shade_bounce(scene, rp.shade, pv, prim, tr.closest, pr, L);
                }
            }
        }
        __syncwarp();
        // ---- path ended: commit it; compat lanes go on with the pixel's next sample
        if (done && !cont) {
            npaths++;
            if (RNG == 0) {
                acc = acc + L;
                L = v3(0.f, 0.f, 0.f);
                sample++;
                if (sample < rp.chunk_start[chunk + 1u]) {
                    float ju = xr.next(), jv = xr.next();
                    start_path_xy(cam, fx, fy, ju, jv, pv);
                    cont = true;
                } else {
                    film.rngA[pixel] = make_uint4(xr.st.v0, xr.st.v1, xr.st.v2, xr.st.v3);
                    film.rngB[pixel] = make_uint2(xr.st.v4, xr.st.d);
                    film.accum[pixel] = make_float4(acc.x, acc.y, acc.z, 0.f);
                    if (chunk + 1u < n_chunks) st_progress(ws.progress + pixel, chunk + 1u);   // release: hand the pixel to whoever holds its next chunk
                    alive = false;
                }
            } else {
                if (L.x != 0.f || L.y != 0.f || L.z != 0.f) {
                    float* fa = reinterpret_cast<float*>(film.accum + pixel);
                    atomicAdd(fa + 0, L.x); atomicAdd(fa + 1, L.y); atomicAdd(fa + 2, L.z);
                }
                alive = false;
            }
        }
        __syncwarp();
        if (done && cont) restart();
    }
    nseg = __reduce_add_sync(kFull, nseg);
    npaths = __reduce_add_sync(kFull, npaths);
    if (lane == 0) {
        if (nseg) atomicAdd(&ws.counters[CTR_SEGMENTS], (unsigned long long)nseg);
        if (npaths) atomicAdd(&ws.counters[CTR_PATHS], (unsigned long long)npaths);
    }
}

// ---------------------------------------------------------------------------------------------
// persistent path loop for TINY scenes (pipeline 6; the default when the scene has a flat unit list, flat.h): the work
// items, the chunked pixel hand-off, shading and regeneration of k_path_persist, but the closest hit comes from
// flat_closest() -- one straight-line pass over <= 20 planar units whose constants are kernel-parameter (constant-bank)
// operands, then one exact quad / triangle-pair test.  Every lane finishes a segment per iteration, so there is no
// traversal state, no stack, no waiting list: all 32 lanes shade together (k_path_persist: 21 of 32 per iteration on the
// Cornell box, and its triangle loop ran at lane efficiency 0.38).  Same hits as the BVH8 path bit for bit (flat.h).
// ---------------------------------------------------------------------------------------------
template <int RNG, int BLOCK, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) k_path_flat(WaveState ws, SceneView scene, const __grid_constant__ FlatScene fs, FilmView film,
                                                         CameraView cam, RenderParams rp) {
    constexpr unsigned kFull = 0xFFFFFFFFu;
    const int lane = (int)(threadIdx.x & 31);
    const uint32_t n_chunks = RNG == 0 ? (uint32_t)rp.n_chunks : 0u;
    const unsigned long long n_items = RNG == 0 ? (unsigned long long)rp.n_owned * n_chunks : rp.total_paths;
    uint32_t nseg = 0, npaths = 0;
    PathVertex pv;
    pv.o = pv.d = pv.beta = v3(0.f, 0.f, 0.f); pv.depth = 0;
    vec3 L = v3(0.f, 0.f, 0.f), acc = v3(0.f, 0.f, 0.f);
    uint32_t pixel = 0, sample = 0, chunk = 0;
    XorwowRng xr; xr.st = Xorwow{0u, 0u, 0u, 0u, 0u, 0u};
    PhiloxRng pr; pr.blk = Philox4{0u, 0u, 0u, 0u}; pr.lane = 0;
    float fx = 0.f, fy = 0.f;
    bool alive = false, pending = false, dry = rp.spp <= 0;
    for (;;) {
        // ---- fetch: every lane without work asks for an item; one atomic per warp
        const bool want = !alive && !pending && !dry;
        const unsigned need = __ballot_sync(kFull, want);
        if (need) {
            unsigned long long base = 0;
            const int leader = __ffs(need) - 1;
            if (lane == leader) base = atomicAdd(&ws.counters[CTR_NEXT_PATH], (unsigned long long)__popc(need));
            base = __shfl_sync(kFull, base, leader);
            if (want) {
                const unsigned long long item = base + (unsigned long long)__popc(need & ((1u << lane) - 1u));
                if (item < n_items) {
                    pixel = ws.slot_pixel[(uint32_t)(item % rp.n_owned)];
                    if (RNG == 0) {
                        chunk = (uint32_t)(item / rp.n_owned);
                        pending = true;
                    } else {
                        sample = (uint32_t)(item / rp.n_owned) + rp.sample_begin;
                        pr.seek(((unsigned long long)sample << 32) | pixel, 0u, rp.seed);
                        float ju = pr.next(), jv = pr.next();
                        start_path_xy(cam, (float)(int)(pixel % (uint32_t)cam.width), (float)(int)(pixel / (uint32_t)cam.width), ju, jv, pv);
                        L = v3(0.f, 0.f, 0.f);
                        alive = true;
                    }
                } else dry = true;
            }
        }
        // ---- compat: start the chunk once its predecessor has been committed (protocol: see k_path_persist)
        if (RNG == 0 && pending) {
            bool ready = chunk == 0u;
            if (!ready) ready = progress_ready(ws.progress + pixel, chunk);
            if (ready) {
                uint4 ra; uint2 rb; float4 f;
                if (chunk == 0u) { ra = film.rngA[pixel]; rb = film.rngB[pixel]; f = film.accum[pixel]; }
                else { ra = __ldcg(film.rngA + pixel); rb = __ldcg(film.rngB + pixel); f = __ldcg(film.accum + pixel); }
                xr.st = Xorwow{ra.x, ra.y, ra.z, ra.w, rb.x, rb.y};
                acc = v3(f.x, f.y, f.z);
                sample = rp.chunk_start[chunk];
                fx = (float)(int)(pixel % (uint32_t)cam.width); fy = (float)(int)(pixel / (uint32_t)cam.width);
                float ju = xr.next(), jv = xr.next();
                start_path_xy(cam, fx, fy, ju, jv, pv);
                L = v3(0.f, 0.f, 0.f);
                pending = false; alive = true;
            }
        }
        if (!__any_sync(kFull, alive || pending)) break;
        // ---- extend: the whole segment
        float t_hit = FLT_MAX; uint32_t prim = kMiss;
        if (alive) flat_closest<false>(fs, pv.o, pv.d, 1e-4f, t_hit, prim, nullptr);
        __syncwarp();
        // ---- shade every live lane
        bool cont = true;
        if (alive) {
            nseg++;
            if (RNG == 0) cont = shade_bounce(scene, rp.shade, pv, prim, t_hit, xr, L);
            else {
                pr.seek(((unsigned long long)sample << 32) | pixel, 1u + (uint32_t)pv.depth, rp.seed);
                cont = shade_bounce(scene, rp.shade, pv, prim, t_hit, pr, L);
            }
        }
        __syncwarp();
        // ---- path ended: commit it; compat lanes go on with the pixel's next sample
        if (alive && !cont) {
            npaths++;
            if (RNG == 0) {
                acc = acc + L;
                L = v3(0.f, 0.f, 0.f);
                sample++;
                if (sample < rp.chunk_start[chunk + 1u]) {
                    float ju = xr.next(), jv = xr.next();
                    start_path_xy(cam, fx, fy, ju, jv, pv);
                } else {
                    film.rngA[pixel] = make_uint4(xr.st.v0, xr.st.v1, xr.st.v2, xr.st.v3);
                    film.rngB[pixel] = make_uint2(xr.st.v4, xr.st.d);
                    film.accum[pixel] = make_float4(acc.x, acc.y, acc.z, 0.f);
                    if (chunk + 1u < n_chunks) st_progress(ws.progress + pixel, chunk + 1u);
                    alive = false;
                }
            } else {
                if (L.x != 0.f || L.y != 0.f || L.z != 0.f) {
                    float* fa = reinterpret_cast<float*>(film.accum + pixel);
                    atomicAdd(fa + 0, L.x); atomicAdd(fa + 1, L.y); atomicAdd(fa + 2, L.z);
                }
                alive = false;
            }
        }
        __syncwarp();
    }
    nseg = __reduce_add_sync(kFull, nseg);
    npaths = __reduce_add_sync(kFull, npaths);
    if (lane == 0) {
        if (nseg) atomicAdd(&ws.counters[CTR_SEGMENTS], (unsigned long long)nseg);
        if (npaths) atomicAdd(&ws.counters[CTR_PATHS], (unsigned long long)npaths);
    }
}

#if defined(PTX_EXPERIMENTAL)   // measured slower than k_path_persist on every scene: not part of the product build
// ---------------------------------------------------------------------------------------------
// CTA wavefront (pipeline 5): ONE launch per frame like k_path_persist, but a path is no longer tied to a
// lane.  A CTA keeps P = BLOCK * PPT paths in SHARED memory (SoA, 4-byte columns) and runs them through four
// phases per iteration, each phase over a dense list of work so that every warp executes it with (nearly) all
// 32 lanes:
//   A  node visits   every traversing path visits ONE BVH8 node (thread t always owns paths t, t + BLOCK, ...:
//                    the traversal cursor and the stack stay in its registers / local memory); the triangles
//                    the node hit become (path, triangle) TASKS in a shared queue; a path with nothing left to
//                    visit goes on the `done` list
//   B  triangle tests one task per lane, result merged with a 64-bit shared atomicMin on { t bits, ~slot }
//   C  shading        one finished path per lane (emission, Russian roulette, albedo, cosine sample)
//   D  regeneration   one ended path per lane: next sample of the pixel / commit + hand-off / fetch of the next
//                    work item (same items, same chunked hand-off protocol as k_path_persist)
// k_path_persist ran the same four pieces per LANE and paid for the divergence: on C2 the triangle loop ran
// 5.0 rounds per iteration for 1.9 tests per lane (lane efficiency 0.33), shading and regeneration at 0.35 /
// 0.32 (ncu r1i).  Here only the last warp of each list is partly filled.
// Hit-distance ties between two triangles resolve towards the higher slot (k_path_persist: the first tested);
// everything else -- arithmetic, draw order, items -- is unchanged, so images match pipelines 0-4 bit for bit
// except at such ties.
// ---------------------------------------------------------------------------------------------
enum : uint32_t { PS_FREE = 0, PS_PENDING = 1, PS_NEW = 2, PS_TRAV = 3, PS_ENDED = 4, PS_DRY = 5 };

template <int P, int CAP>
struct CtaPaths {
    float o[3][P], d[3][P];
    unsigned long long key[P];        // { closest t (float bits), ~triangle slot }: smaller is closer; low word 0 = no hit
    uint32_t tgx[P];                  // base slot of the triangle group the path's tasks refer to
    float beta[3][P], L[3][P], acc[3][P];
    uint32_t depth[P], pixel[P], pxy[P], sample[P], chunk[P], status[P];
    uint32_t rng[6][P];
    uint16_t tasks[CAP];              // (path << 5) | bit within the triangle group; 0xFFFF = hole
    uint16_t done[P], regen[P];
    uint32_t cnt[2][4];               // [iteration parity][tasks, done, regen, -]
};

constexpr unsigned long long kKeyInit = 0x7F7FFFFF00000000ull;     // FLT_MAX, no hit

template <int RNG, int BLOCK, int PPT, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) k_path_cta(WaveState ws, SceneView scene, FilmView film, CameraView cam, RenderParams rp) {
    constexpr unsigned kFull = 0xFFFFFFFFu;
    constexpr int P = BLOCK * PPT;
    constexpr int CAP = 4 * P;
    static_assert(P <= 2048, "task encoding holds 11 path bits");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    CtaPaths<P, CAP>& S = *reinterpret_cast<CtaPaths<P, CAP>*>(smem_raw);
    const int tid = (int)threadIdx.x, lane = tid & 31;
    const unsigned lt = (1u << lane) - 1u;
    const uint32_t n_chunks = RNG == 0 ? (uint32_t)rp.n_chunks : 0u;
    const unsigned long long n_items = RNG == 0 ? (unsigned long long)rp.n_owned * n_chunks : rp.total_paths;
    uint32_t nseg = 0, npaths = 0;
    Traversal tr[PPT];
    TravStack<24> stk[PPT];
#pragma unroll
    for (int i = 0; i < PPT; i++) {
        S.status[tid + i * BLOCK] = rp.spp > 0 ? PS_FREE : PS_DRY;
        tr[i].begin(v3(0.f, 0.f, 0.f), FLT_MAX);
    }
    if (tid < 8) (&S.cnt[0][0])[tid] = 0u;
    __syncthreads();

    for (uint32_t it = 0;; it++) {
        uint32_t* cnt = S.cnt[it & 1u];
        // ================= A: one node visit per traversing path =================
        bool any_live = false;
#pragma unroll
        for (int i = 0; i < PPT; i++) {
            const int p = tid + i * BLOCK;
            const uint32_t st = S.status[p];
            any_live |= st != PS_DRY;
            uint32_t tg_x = 0, tg_y = 0;
            bool fin = false;
            vec3 o = v3(0.f, 0.f, 0.f), d = o;
            if (st == PS_NEW || st == PS_TRAV) {
                o = v3(S.o[0][p], S.o[1][p], S.o[2][p]);
                d = v3(S.d[0][p], S.d[1][p], S.d[2][p]);
                if (st == PS_NEW) { tr[i].begin(d, FLT_MAX); S.status[p] = PS_TRAV; }
                tr[i].closest = __uint_as_float(reinterpret_cast<const uint32_t*>(&S.key[p])[1]);
                tr[i].template visit<false>(scene, stk[i], o, d, 1e-4f, tg_x, tg_y, nullptr);
                fin = tr[i].next_group(stk[i]);
            }
            __syncwarp();
            // triangle tasks: warp prefix sum of the per-lane counts, one shared atomic per warp
            const uint32_t n = (uint32_t)__popc(tg_y);
            uint32_t incl = n;
#pragma unroll
            for (int k = 1; k < 32; k <<= 1) {
                uint32_t v = __shfl_up_sync(kFull, incl, k);
                if (lane >= k) incl += v;
            }
            const uint32_t total = __shfl_sync(kFull, incl, 31);
            if (total) {
                uint32_t base = 0;
                if (lane == 31) base = atomicAdd(&cnt[0], total);
                base = __shfl_sync(kFull, base, 31);
                uint32_t pos = base + incl - n;
                if (n) {
                    if (pos + n <= (uint32_t)CAP) {
                        S.tgx[p] = tg_x;
                        const uint32_t tag = (uint32_t)p << 5;
                        while (tg_y) {
                            uint32_t bit = hi_bit(tg_y);
                            tg_y &= ~(1u << bit);
                            S.tasks[pos++] = (uint16_t)(tag | bit);
                        }
                    } else {
                        // queue full (a node can hit 24 triangles): test them here, mark the reserved slots as holes
                        for (; pos < (uint32_t)CAP; pos++) S.tasks[pos] = (uint16_t)0xFFFFu;
                        unsigned long long k = S.key[p];
                        float c = __uint_as_float((uint32_t)(k >> 32));
                        uint32_t best = ~(uint32_t)k;
                        while (tg_y) {
                            uint32_t bit = hi_bit(tg_y);
                            tg_y &= ~(1u << bit);
                            uint32_t ti = tg_x + bit;
                            const f32x4* tp = scene.tris + (size_t)ti * kTriQuads;
                            f32x4 a = ldq(tp + 0), b = ldq(tp + 1), e = ldq(tp + 2);
                            if (intersect_tri(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(e.x, e.y, e.z), o, d, 1e-4f, c)) best = ti;
                        }
                        S.key[p] = ((unsigned long long)__float_as_uint(c) << 32) | (unsigned long long)(~best);
                    }
                }
            }
            const unsigned mfin = __ballot_sync(kFull, fin);
            if (mfin) {
                uint32_t base = 0;
                const int leader = __ffs(mfin) - 1;
                if (lane == leader) base = atomicAdd(&cnt[1], (uint32_t)__popc(mfin));
                base = __shfl_sync(kFull, base, leader);
                if (fin) S.done[base + (uint32_t)__popc(mfin & lt)] = (uint16_t)p;
            }
            const bool idle = st == PS_FREE || st == PS_PENDING;
            const unsigned midle = __ballot_sync(kFull, idle);
            if (midle) {
                uint32_t base = 0;
                const int leader = __ffs(midle) - 1;
                if (lane == leader) base = atomicAdd(&cnt[2], (uint32_t)__popc(midle));
                base = __shfl_sync(kFull, base, leader);
                if (idle) S.regen[base + (uint32_t)__popc(midle & lt)] = (uint16_t)p;
            }
        }
        if (!__syncthreads_or(any_live ? 1 : 0)) break;

        // ================= B: one triangle test per task =================
        {
            const uint32_t nt = min(cnt[0], (uint32_t)CAP);
            for (uint32_t k = (uint32_t)tid; k < nt; k += BLOCK) {
                const uint32_t tk = S.tasks[k];
                if (tk == 0xFFFFu) continue;
                const int p = (int)(tk >> 5);
                const uint32_t ti = S.tgx[p] + (tk & 31u);
                const f32x4* tp = scene.tris + (size_t)ti * kTriQuads;
                f32x4 a = ldq(tp + 0), b = ldq(tp + 1), e = ldq(tp + 2);
                vec3 o = v3(S.o[0][p], S.o[1][p], S.o[2][p]);
                vec3 d = v3(S.d[0][p], S.d[1][p], S.d[2][p]);
                float c = __uint_as_float(reinterpret_cast<const uint32_t*>(&S.key[p])[1]);
                if (intersect_tri(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(e.x, e.y, e.z), o, d, 1e-4f, c))
                    atomicMin(&S.key[p], ((unsigned long long)__float_as_uint(c) << 32) | (unsigned long long)(~ti));
            }
        }
        __syncthreads();

        // ================= C: shade the paths whose traversal is complete =================
        {
            const uint32_t nd = cnt[1];
            for (uint32_t k0 = (uint32_t)(tid & ~31); k0 < nd; k0 += BLOCK) {
                const uint32_t k = k0 + (uint32_t)lane;
                const bool act = k < nd;
                bool ended = false;
                int p = 0;
                if (act) {
                    p = (int)S.done[k];
                    const unsigned long long key = S.key[p];
                    const float t = __uint_as_float((uint32_t)(key >> 32));
                    const uint32_t best = ~(uint32_t)key;
                    const uint32_t prim = best == kMiss ? kMiss : __float_as_uint(ldq(scene.tris + (size_t)best * kTriQuads).w);
                    PathVertex pv;
                    pv.o = v3(S.o[0][p], S.o[1][p], S.o[2][p]);
                    pv.d = v3(S.d[0][p], S.d[1][p], S.d[2][p]);
                    pv.beta = v3(S.beta[0][p], S.beta[1][p], S.beta[2][p]);
                    pv.depth = (int)S.depth[p];
                    vec3 L = v3(S.L[0][p], S.L[1][p], S.L[2][p]);
                    nseg++;
                    bool cont;
                    if (RNG == 0) {
                        XorwowRng xr;
                        xr.st = Xorwow{S.rng[0][p], S.rng[1][p], S.rng[2][p], S.rng[3][p], S.rng[4][p], S.rng[5][p]};
                        cont = shade_bounce(scene, rp.shade, pv, prim, t, xr, L);
                        if (prim != kMiss) {
                            S.rng[0][p] = xr.st.v0; S.rng[1][p] = xr.st.v1; S.rng[2][p] = xr.st.v2;
                            S.rng[3][p] = xr.st.v3; S.rng[4][p] = xr.st.v4; S.rng[5][p] = xr.st.d;
                        }
                    } else {
                        PhiloxRng pr;
                        pr.seek(((unsigned long long)S.sample[p] << 32) | S.pixel[p], 1u + (uint32_t)pv.depth, rp.seed);
                        cont = shade_bounce(scene, rp.shade, pv, prim, t, pr, L);
                    }
                    S.L[0][p] = L.x; S.L[1][p] = L.y; S.L[2][p] = L.z;
                    if (cont) {
                        S.o[0][p] = pv.o.x; S.o[1][p] = pv.o.y; S.o[2][p] = pv.o.z;
                        S.d[0][p] = pv.d.x; S.d[1][p] = pv.d.y; S.d[2][p] = pv.d.z;
                        S.beta[0][p] = pv.beta.x; S.beta[1][p] = pv.beta.y; S.beta[2][p] = pv.beta.z;
                        S.depth[p] = (uint32_t)pv.depth;
                        S.key[p] = kKeyInit;
                        S.status[p] = PS_NEW;
                    } else {
                        S.status[p] = PS_ENDED;
                        ended = true;
                    }
                }
                const unsigned mend = __ballot_sync(kFull, ended);
                if (mend) {
                    uint32_t base = 0;
                    const int leader = __ffs(mend) - 1;
                    if (lane == leader) base = atomicAdd(&cnt[2], (uint32_t)__popc(mend));
                    base = __shfl_sync(kFull, base, leader);
                    if (ended) S.regen[base + (uint32_t)__popc(mend & lt)] = (uint16_t)p;
                }
            }
        }
        __syncthreads();

        // ================= D: ended / idle paths: next sample, commit + hand-off, next work item =================
        {
            const uint32_t nr = cnt[2];
            if (tid < 4) S.cnt[(it + 1u) & 1u][tid] = 0u;      // the other counter set is idle during this phase
            for (uint32_t k0 = (uint32_t)(tid & ~31); k0 < nr; k0 += BLOCK) {
                const uint32_t k = k0 + (uint32_t)lane;
                const bool act = k < nr;
                const int p = act ? (int)S.regen[k] : 0;
                uint32_t st = act ? S.status[p] : PS_DRY;
                bool want = st == PS_FREE;
                if (st == PS_ENDED) {
                    npaths++;
                    const uint32_t pixel = S.pixel[p];
                    const vec3 L = v3(S.L[0][p], S.L[1][p], S.L[2][p]);
                    if (RNG == 0) {
                        const uint32_t chunk = S.chunk[p];
                        const uint32_t sample = S.sample[p] + 1u;
                        const vec3 acc = v3(S.acc[0][p], S.acc[1][p], S.acc[2][p]) + L;
                        XorwowRng xr;
                        xr.st = Xorwow{S.rng[0][p], S.rng[1][p], S.rng[2][p], S.rng[3][p], S.rng[4][p], S.rng[5][p]};
                        if (sample < rp.chunk_start[chunk + 1u]) {
                            const uint32_t pxy = S.pxy[p];
                            float ju = xr.next(), jv = xr.next();
                            PathVertex pv;
                            start_path_xy(cam, (float)(int)(pxy & 0xFFFFu), (float)(int)(pxy >> 16), ju, jv, pv);
                            S.o[0][p] = pv.o.x; S.o[1][p] = pv.o.y; S.o[2][p] = pv.o.z;
                            S.d[0][p] = pv.d.x; S.d[1][p] = pv.d.y; S.d[2][p] = pv.d.z;
                            S.beta[0][p] = 1.0f; S.beta[1][p] = 1.0f; S.beta[2][p] = 1.0f;
                            S.depth[p] = 0u;
                            S.L[0][p] = 0.f; S.L[1][p] = 0.f; S.L[2][p] = 0.f;
                            S.acc[0][p] = acc.x; S.acc[1][p] = acc.y; S.acc[2][p] = acc.z;
                            S.sample[p] = sample;
                            S.rng[0][p] = xr.st.v0; S.rng[1][p] = xr.st.v1; S.rng[2][p] = xr.st.v2;
                            S.rng[3][p] = xr.st.v3; S.rng[4][p] = xr.st.v4; S.rng[5][p] = xr.st.d;
                            S.key[p] = kKeyInit;
                            st = PS_NEW;
                        } else {
                            film.rngA[pixel] = make_uint4(xr.st.v0, xr.st.v1, xr.st.v2, xr.st.v3);
                            film.rngB[pixel] = make_uint2(xr.st.v4, xr.st.d);
                            film.accum[pixel] = make_float4(acc.x, acc.y, acc.z, 0.f);
                            if (chunk + 1u < n_chunks)                                // release: hand the pixel to whoever holds its next chunk
                                asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(ws.progress + pixel), "r"(chunk + 1u) : "memory");
                            want = true;
                        }
                    } else {
                        if (L.x != 0.f || L.y != 0.f || L.z != 0.f) {
                            float* fa = reinterpret_cast<float*>(film.accum + pixel);
                            atomicAdd(fa + 0, L.x); atomicAdd(fa + 1, L.y); atomicAdd(fa + 2, L.z);
                        }
                        want = true;
                    }
                }
                // ---- fetch: one global atomic per warp
                const unsigned need = __ballot_sync(kFull, want);
                if (need) {
                    unsigned long long base = 0;
                    const int leader = __ffs(need) - 1;
                    if (lane == leader) base = atomicAdd(&ws.counters[CTR_NEXT_PATH], (unsigned long long)__popc(need));
                    base = __shfl_sync(kFull, base, leader);
                    if (want) {
                        const unsigned long long item = base + (unsigned long long)__popc(need & lt);
                        if (item < n_items) {
                            const uint32_t pixel = ws.slot_pixel[(uint32_t)(item % rp.n_owned)];
                            const uint32_t px = pixel % (uint32_t)cam.width, py = pixel / (uint32_t)cam.width;
                            S.pixel[p] = pixel;
                            S.pxy[p] = px | (py << 16);
                            if (RNG == 0) {
                                S.chunk[p] = (uint32_t)(item / rp.n_owned);
                                st = PS_PENDING;
                            } else {
                                const uint32_t sample = (uint32_t)(item / rp.n_owned) + rp.sample_begin;
                                PhiloxRng pr;
                                pr.seek(((unsigned long long)sample << 32) | pixel, 0u, rp.seed);
                                float ju = pr.next(), jv = pr.next();
                                PathVertex pv;
                                start_path_xy(cam, (float)(int)px, (float)(int)py, ju, jv, pv);
                                S.o[0][p] = pv.o.x; S.o[1][p] = pv.o.y; S.o[2][p] = pv.o.z;
                                S.d[0][p] = pv.d.x; S.d[1][p] = pv.d.y; S.d[2][p] = pv.d.z;
                                S.beta[0][p] = 1.0f; S.beta[1][p] = 1.0f; S.beta[2][p] = 1.0f;
                                S.depth[p] = 0u;
                                S.L[0][p] = 0.f; S.L[1][p] = 0.f; S.L[2][p] = 0.f;
                                S.sample[p] = sample;
                                S.key[p] = kKeyInit;
                                st = PS_NEW;
                            }
                        } else st = PS_DRY;
                    }
                }
                // ---- compat: start the chunk once its predecessor has been committed (see k_path_persist)
                if (RNG == 0 && st == PS_PENDING) {
                    const uint32_t pixel = S.pixel[p], chunk = S.chunk[p];
                    bool ready = chunk == 0u;
                    if (!ready) {
                        uint32_t seen;
                        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(ws.progress + pixel) : "memory");
                        ready = seen == chunk;
                    }
                    if (ready) {
                        uint4 ra; uint2 rb; float4 f;
                        if (chunk == 0u) { ra = film.rngA[pixel]; rb = film.rngB[pixel]; f = film.accum[pixel]; }
                        else { ra = __ldcg(film.rngA + pixel); rb = __ldcg(film.rngB + pixel); f = __ldcg(film.accum + pixel); }
                        XorwowRng xr;
                        xr.st = Xorwow{ra.x, ra.y, ra.z, ra.w, rb.x, rb.y};
                        const uint32_t pxy = S.pxy[p];
                        float ju = xr.next(), jv = xr.next();
                        PathVertex pv;
                        start_path_xy(cam, (float)(int)(pxy & 0xFFFFu), (float)(int)(pxy >> 16), ju, jv, pv);
                        S.o[0][p] = pv.o.x; S.o[1][p] = pv.o.y; S.o[2][p] = pv.o.z;
                        S.d[0][p] = pv.d.x; S.d[1][p] = pv.d.y; S.d[2][p] = pv.d.z;
                        S.beta[0][p] = 1.0f; S.beta[1][p] = 1.0f; S.beta[2][p] = 1.0f;
                        S.depth[p] = 0u;
                        S.L[0][p] = 0.f; S.L[1][p] = 0.f; S.L[2][p] = 0.f;
                        S.acc[0][p] = f.x; S.acc[1][p] = f.y; S.acc[2][p] = f.z;
                        S.sample[p] = rp.chunk_start[chunk];
                        S.rng[0][p] = xr.st.v0; S.rng[1][p] = xr.st.v1; S.rng[2][p] = xr.st.v2;
                        S.rng[3][p] = xr.st.v3; S.rng[4][p] = xr.st.v4; S.rng[5][p] = xr.st.d;
                        S.key[p] = kKeyInit;
                        st = PS_NEW;
                    }
                }
                if (act) S.status[p] = st;
            }
        }
        __syncthreads();
    }
    nseg = __reduce_add_sync(kFull, nseg);
    npaths = __reduce_add_sync(kFull, npaths);
    if (lane == 0) {
        if (nseg) atomicAdd(&ws.counters[CTR_SEGMENTS], (unsigned long long)nseg);
        if (npaths) atomicAdd(&ws.counters[CTR_PATHS], (unsigned long long)npaths);
    }
}

#endif  // PTX_EXPERIMENTAL

// ---------------------------------------------------------------------------------------------
// Work ORDER of the persistent kernels, compat RNG (longest-expected pixels first).  A pixel's samples are one sequential chain, and
// the work items are handed out in slot order: a pixel with long paths (the interior of the box) whose first chunk is fetched late
// finishes after everything else and stretches the frame's tail (one rank of eight on C2: 20.7 ms for 16.9 ms of work).  This probe
// traces kProbePaths paths per owned pixel with a counter RNG of its own -- it never touches the film's XORWOW streams -- and counts
// their segments; the slots are then sorted by that cost, descending.  Only the ORDER of the work changes: the film is identical.
// ---------------------------------------------------------------------------------------------
constexpr int kProbePaths = 2;
template <bool FLAT>
__global__ void __launch_bounds__(kBlock) k_pixel_cost(SceneView scene, const __grid_constant__ FlatScene fs, CameraView cam, ShadeParams sp,
                                                       const uint32_t* __restrict__ slot_base, uint32_t n_owned, uint32_t* keys, uint32_t* vals) {
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= n_owned) return;
    const uint32_t pixel = slot_base[i];
    const float fx = (float)(int)(pixel % (uint32_t)cam.width), fy = (float)(int)(pixel / (uint32_t)cam.width);
    uint32_t nseg = 0;
    for (int k = 0; k < kProbePaths; k++) {
        PhiloxRng pr;
        pr.seek(((unsigned long long)(0xC0570000u + (uint32_t)k) << 32) | pixel, 0u, 0x9E3779B97F4A7C15ull);
        PathVertex pv;
        float ju = pr.next(), jv = pr.next();
        start_path_xy(cam, fx, fy, ju, jv, pv);
        vec3 L = v3(0.f, 0.f, 0.f);
        bool cont = true;
        while (cont) {
            float t; uint32_t prim;
            if (FLAT) flat_closest<false>(fs, pv.o, pv.d, 1e-4f, t, prim, nullptr);
            else {
                uint32_t tri = trace_closest<24, false>(scene, pv.o, pv.d, 1e-4f, FLT_MAX, t, nullptr);
                prim = tri == kMiss ? kMiss : __float_as_uint(ldq(scene.tris + (size_t)tri * kTriQuads).w);
            }
            nseg++;
            pr.seek(((unsigned long long)(0xC0570000u + (uint32_t)k) << 32) | pixel, 1u + (uint32_t)pv.depth, 0x9E3779B97F4A7C15ull);
            cont = shade_bounce(scene, sp, pv, prim, t, pr, L);
        }
    }
    keys[i] = 255u - min(nseg, 255u);          // ascending sort of this key = descending cost; stable: equal costs keep slot order
    vals[i] = pixel;
}

// ---------------------------------------------------------------------------------------------
// Queue sorting for the wavefront pipeline (north_star subsystem 3: ray reordering + material-sorted shading).
//   ray key      : 15-bit Morton code of the ray origin's cell (32^3 over the scene bounds) << 3 | direction octant -- rays that start
//                  in the same region and head the same way visit the same nodes and triangle records (L1 / L2 hits in extend)
//   material key : material index of the primitive the ray hit (misses after every material) -- lanes of a warp take the same
//                  branch of shade_bounce (emitter / mirror / diffuse / miss)
// Entries past the live length and dead slots get the largest keys, so a stable sort keeps them behind the live ones.
// The order of the queue cannot change an image: a slot's path, RNG stream and (compat) pixel are its own.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t spread5(uint32_t v) {          // 5 bits -> every third bit
    v = (v | (v << 8)) & 0x0000F00Fu; v = (v | (v << 4)) & 0x000C30C3u; v = (v | (v << 2)) & 0x00249249u;
    return v;
}
__global__ void __launch_bounds__(kBlock) k_queue_keys(WaveState ws, SceneView scene, int q, uint32_t n_bound, int by_material, float3 lo, float3 inv_ext,
                                                       uint32_t* keys) {
    uint32_t k = blockIdx.x * kBlock + threadIdx.x;
    if (k >= n_bound) return;
    uint32_t qlen = (uint32_t)ws.counters[CTR_QLEN];
    uint32_t key = 0x7FFFFFFFu;
    if (k < qlen) {
        uint32_t slot = ws.queue[q][k];
        float4 b = ws.B[slot];
        key = 0x7FFFFFFEu;
        if (!(__float_as_uint(b.z) & kDeadFlag)) {
            if (by_material) {
                uint32_t prim = __float_as_uint(ws.C[slot].w);
                key = prim == kMiss ? 0x0000FFFFu : (f2u(ldq(scene.prim_shade + prim).w) & 0xFFFFu);
            } else {
                float4 a = ws.A[slot];
                uint32_t cx = (uint32_t)fminf(fmaxf((a.x - lo.x) * inv_ext.x, 0.0f), 31.0f);
                uint32_t cy = (uint32_t)fminf(fmaxf((a.y - lo.y) * inv_ext.y, 0.0f), 31.0f);
                uint32_t cz = (uint32_t)fminf(fmaxf((a.z - lo.z) * inv_ext.z, 0.0f), 31.0f);
                uint32_t oct = (a.w < 0.0f ? 4u : 0u) | (b.x < 0.0f ? 2u : 0u) | (b.y < 0.0f ? 1u : 0u);
                key = ((spread5(cx) | (spread5(cy) << 1) | (spread5(cz) << 2)) << 3) | oct;
            }
        }
    }
    keys[k] = key;
}

// ---------------------------------------------------------------------------------------------
// compact: stable removal of dead slots from the queue.  count -> scan -> scatter -> commit.
// Each block owns 1024 consecutive queue entries; order is preserved so state accesses stay coalesced.
// ---------------------------------------------------------------------------------------------
constexpr int kCompactChunk = 1024;

__device__ __forceinline__ bool slot_live(const WaveState& ws, uint32_t slot) {
    return !(__float_as_uint(ws.B[slot].z) & kDeadFlag);
}

__global__ void __launch_bounds__(kBlock) k_compact_count(WaveState ws, int q, uint32_t* block_counts) {
    uint32_t qlen = (uint32_t)ws.counters[CTR_QLEN];
    __shared__ uint32_t total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    uint32_t mine = 0;
    for (int r = 0; r < kCompactChunk / kBlock; r++) {
        uint32_t k = blockIdx.x * kCompactChunk + r * kBlock + threadIdx.x;
        if (k < qlen && slot_live(ws, ws.queue[q][k])) mine++;
    }
    mine = __reduce_add_sync(0xFFFFFFFFu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&total, mine);
    __syncthreads();
    if (threadIdx.x == 0) block_counts[blockIdx.x] = total;
}

// single block: exclusive scan of block_counts[0..nb) in place; block_counts[nb] = total
__global__ void __launch_bounds__(1024) k_compact_scan(uint32_t* block_counts, uint32_t nb) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nb; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint32_t v = i < nb ? block_counts[i] : 0u;
        uint32_t incl = v;
        for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o); if ((threadIdx.x & 31) >= o) incl += n; }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sums[threadIdx.x], wi = w;
            for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xFFFFFFFFu, wi, o); if ((int)threadIdx.x >= o) wi += n; }
            warp_sums[threadIdx.x] = wi - w;      // exclusive
        }
        __syncthreads();
        uint32_t excl = carry + warp_sums[threadIdx.x >> 5] + incl - v;
        if (i < nb) block_counts[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) block_counts[nb] = carry;
}

__global__ void __launch_bounds__(kBlock) k_compact_scatter(WaveState ws, int q, const uint32_t* block_offsets) {
    uint32_t qlen = (uint32_t)ws.counters[CTR_QLEN];
    __shared__ uint32_t warp_counts[kBlock / 32];
    __shared__ uint32_t running;
    if (threadIdx.x == 0) running = block_offsets[blockIdx.x];
    __syncthreads();
    const uint32_t* src = ws.queue[q];
    uint32_t* dst = ws.queue[q ^ 1];
    for (int r = 0; r < kCompactChunk / kBlock; r++) {
        uint32_t k = blockIdx.x * kCompactChunk + r * kBlock + threadIdx.x;
        uint32_t slot = k < qlen ? src[k] : 0u;
        bool keep = k < qlen && slot_live(ws, slot);
        unsigned m = __ballot_sync(0xFFFFFFFFu, keep);
        int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        if (lane == 0) warp_counts[warp] = __popc(m);
        __syncthreads();
        uint32_t before = running;
        for (int w = 0; w < warp; w++) before += warp_counts[w];
        if (keep) dst[before + __popc(m & ((1u << lane) - 1u))] = slot;
        __syncthreads();
        if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < kBlock / 32; w++) t += warp_counts[w]; running += t; }
        __syncthreads();
    }
}

__global__ void k_compact_commit(WaveState ws, const uint32_t* block_offsets, uint32_t nb) {
    ws.counters[CTR_QLEN] = block_offsets[nb];
}

// ---------------------------------------------------------------------------------------------
// resolve / film read-back (integrator.h:393-407)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_resolve(FilmView film, float inv_spp, uint8_t* rgb8, int flip_y) {
    uint32_t p = blockIdx.x * kBlock + threadIdx.x;
    uint32_t npix = (uint32_t)film.width * (uint32_t)film.height;
    if (p >= npix) return;
    float4 a = film.accum[p];
    uint32_t x = p % (uint32_t)film.width, y = p / (uint32_t)film.width;
    uint32_t o = flip_y ? ((uint32_t)film.height - 1u - y) * (uint32_t)film.width + x : p;
    rgb8[3 * (size_t)o + 0] = tonemap_channel(a.x * inv_spp);
    rgb8[3 * (size_t)o + 1] = tonemap_channel(a.y * inv_spp);
    rgb8[3 * (size_t)o + 2] = tonemap_channel(a.z * inv_spp);
}

__global__ void __launch_bounds__(kBlock) k_film_scale(FilmView film, float inv_spp, float* linear_rgb) {
    uint32_t p = blockIdx.x * kBlock + threadIdx.x;
    uint32_t npix = (uint32_t)film.width * (uint32_t)film.height;
    if (p >= npix) return;
    float4 a = film.accum[p];
    linear_rgb[3 * (size_t)p + 0] = a.x * inv_spp;
    linear_rgb[3 * (size_t)p + 1] = a.y * inv_spp;
    linear_rgb[3 * (size_t)p + 2] = a.z * inv_spp;
}

// ---------------------------------------------------------------------------------------------
// film combine, tile sharding (ptx_film_combine): every rank packs the running sums of the pixels it owns into its slice
// of the gather buffer, one ncclAllGather moves 1 / G of the film per rank, and every rank unpacks all slices into its film.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_film_pack_owned(FilmView film, const uint32_t* __restrict__ slot_pixel, uint32_t n_owned, float4* dst) {
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i < n_owned) dst[i] = film.accum[slot_pixel[i]];
}
// inverse of k_slot_pixels for every rank at once: pixel -> (owner, local slot)
__global__ void __launch_bounds__(kBlock) k_film_unpack_all(FilmView film, const float4* __restrict__ gathered, uint32_t per_rank, int world) {
    uint32_t p = blockIdx.x * kBlock + threadIdx.x;
    uint32_t npix = (uint32_t)film.width * (uint32_t)film.height;
    if (p >= npix) return;
    const uint32_t band_px = (uint32_t)kBandRows * (uint32_t)film.width;
    const uint32_t band = p / band_px, within = p - band * band_px;
    const uint32_t owner = band % (uint32_t)world, local_band = band / (uint32_t)world;
    film.accum[p] = gathered[(size_t)owner * per_rank + (size_t)local_band * band_px + within];
}

// ---------------------------------------------------------------------------------------------
// queries for parity tests
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_primary_ids(SceneView scene, CameraView cam, int w, int h, int32_t* prim, float* tout) {
    uint32_t p = blockIdx.x * kBlock + threadIdx.x;
    if (p >= (uint32_t)w * (uint32_t)h) return;
    int x = (int)(p % (uint32_t)w), y = (int)(p / (uint32_t)w);
    float u = ((float)x + 0.5f) / (float)w, v = ((float)y + 0.5f) / (float)h;
    vec3 o, d; camera_ray(cam, u, v, o, d);
    float t;
    uint32_t tri = trace_closest<24, false>(scene, o, d, 1e-4f, FLT_MAX, t, nullptr);
    prim[p] = tri == kMiss ? -1 : (int32_t)__float_as_uint(ldq(scene.tris + (size_t)tri * kTriQuads).w);
    tout[p] = tri == kMiss ? -1.0f : t;
}

__global__ void __launch_bounds__(kBlock) k_trace_rays(SceneView scene, const float* __restrict__ rays, size_t n, float t_min, int32_t* prim,
                                                       float* tout) {
    size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n) return;
    vec3 o = v3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
    vec3 d = unit(v3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]));   // Ray ctor normalises (core/ray.h:9-12)
    float t;
    uint32_t tri = trace_closest<24, false>(scene, o, d, t_min, FLT_MAX, t, nullptr);
    prim[i] = tri == kMiss ? -1 : (int32_t)__float_as_uint(ldq(scene.tris + (size_t)tri * kTriQuads).w);
    tout[i] = tri == kMiss ? -1.0f : t;
}

// two-level scenes: the same queries through the TLAS / BLAS traversal; ids in the flattened numbering
__global__ void __launch_bounds__(kBlock) k_primary_ids_inst(SceneView scene, CameraView cam, int w, int h, int32_t* prim, float* tout) {
    uint32_t p = blockIdx.x * kBlock + threadIdx.x;
    if (p >= (uint32_t)w * (uint32_t)h) return;
    int x = (int)(p % (uint32_t)w), y = (int)(p / (uint32_t)w);
    float u = ((float)x + 0.5f) / (float)w, v = ((float)y + 0.5f) / (float)h;
    vec3 o, d; camera_ray(cam, u, v, o, d);
    float t; uint32_t id;
    uint32_t tri = trace_closest_instanced<24>(scene, o, d, 1e-4f, FLT_MAX, t, id);
    prim[p] = tri == kMiss ? -1 : (int32_t)id;
    tout[p] = tri == kMiss ? -1.0f : t;
}
__global__ void __launch_bounds__(kBlock) k_trace_rays_inst(SceneView scene, const float* __restrict__ rays, size_t n, float t_min, int32_t* prim, float* tout) {
    size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n) return;
    vec3 o = v3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
    vec3 d = unit(v3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]));
    float t; uint32_t id;
    uint32_t tri = trace_closest_instanced<24>(scene, o, d, t_min, FLT_MAX, t, id);
    prim[i] = tri == kMiss ? -1 : (int32_t)id;
    tout[i] = tri == kMiss ? -1.0f : t;
}
// a BLAS copied behind other trees in the shared node / triangle arrays: shift its child and triangle bases
__global__ void __launch_bounds__(kBlock) k_offset_nodes(u32x4* nodes, uint32_t n_nodes, uint32_t node_off, uint32_t tri_off) {
    uint32_t i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= n_nodes) return;
    u32x4 q = nodes[(size_t)i * kNodeQuads + 1];
    q.x += node_off; q.y += tri_off;
    nodes[(size_t)i * kNodeQuads + 1] = q;
}

// Scene::occluded(from, to) (scene.h:132-149): the shadow ray starts 1e-4 along the normalised direction, is normalised
// again by the Ray ctor, and any hit in [1e-4, dist - 1e-4) occludes.
__global__ void __launch_bounds__(kBlock) k_occluded(SceneView scene, const float* __restrict__ segs, size_t n, uint8_t* out) {
    size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n) return;
    vec3 from = v3(segs[6 * i], segs[6 * i + 1], segs[6 * i + 2]);
    vec3 dir = v3(segs[6 * i + 3], segs[6 * i + 4], segs[6 * i + 5]) - from;
    float dist = length(dir);
    dir = div_scalar(dir, dist);
    vec3 o = from + 1e-4f * dir;
    vec3 d = unit(dir);
    out[i] = trace_any<24>(scene, o, d, 1e-4f, dist - 1e-4f) ? 1 : 0;
}

__global__ void __launch_bounds__(kBlock) k_occluded_inst(SceneView scene, const float* __restrict__ segs, size_t n, uint8_t* out) {
    size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n) return;
    vec3 from = v3(segs[6 * i], segs[6 * i + 1], segs[6 * i + 2]);
    vec3 dir = v3(segs[6 * i + 3], segs[6 * i + 4], segs[6 * i + 5]) - from;
    float dist = length(dir);
    dir = div_scalar(dir, dist);
    vec3 o = from + 1e-4f * dir;
    vec3 d = unit(dir);
    out[i] = trace_any_instanced<24>(scene, o, d, 1e-4f, dist - 1e-4f) ? 1 : 0;
}

// render_radiosity (integrator.h:460-504): per pixel, spp jittered primary rays; colour = Le + the radiosity stored for
// the hit primitive (input of this kernel: the solver that produces it is out of scope); sqrt "gamma", clamp to 1.
// Continues and writes back the pixel's XORWOW stream like the reference kernel.
template <bool INST>
__global__ void __launch_bounds__(kBlock) k_visibility(SceneView scene, FilmView film, CameraView cam, int spp, const float* __restrict__ radiosity,
                                                       uint8_t* rgb8) {
    uint32_t p = blockIdx.x * kBlock + threadIdx.x;
    if (p >= (uint32_t)cam.width * (uint32_t)cam.height) return;
    float fx = (float)(int)(p % (uint32_t)cam.width), fy = (float)(int)(p / (uint32_t)cam.width);
    uint4 ra = film.rngA[p]; uint2 rb = film.rngB[p];
    XorwowRng xr; xr.st = Xorwow{ra.x, ra.y, ra.z, ra.w, rb.x, rb.y};
    vec3 color = v3(0.f, 0.f, 0.f);
    for (int s = 0; s < spp; s++) {
        float u = (fx + xr.next()) / (float)cam.width;
        float v = (fy + xr.next()) / (float)cam.height;
        vec3 o, d; camera_ray(cam, u, v, o, d);
        float t;
        uint32_t tri, prim = kMiss, mat = 0;
        if constexpr (INST) {
            InstTraversal it;
            TravStack<24> stk;
            it.begin(o, d, FLT_MAX);
            while (!it.template step<24>(scene, stk, o, d, 1e-4f)) {}
            tri = it.tr.best; t = it.tr.closest;
            if (tri != kMiss) { vec3 n; resolve_instanced_hit(scene, it, prim, n, mat); }
        } else {
            tri = trace_closest<24, false>(scene, o, d, 1e-4f, FLT_MAX, t, nullptr);
            if (tri != kMiss) {
                prim = __float_as_uint(ldq(scene.tris + (size_t)tri * kTriQuads).w);
                mat = f2u(ldq(scene.prim_shade + prim).w);
            }
        }
        if (tri != kMiss) {
            f32x4 ke = ldq(scene.materials + 2 * (size_t)mat + 1);
            color = color + v3(ke.x, ke.y, ke.z);
            if (radiosity) color = color + v3(radiosity[3 * (size_t)prim], radiosity[3 * (size_t)prim + 1], radiosity[3 * (size_t)prim + 2]);
            else color = color + v3(0.f, 0.f, 0.f);
        }
    }
    color = div_scalar(color, (float)spp);
    rgb8[3 * (size_t)p + 0] = (uint8_t)(255.99f * sqrt_exact(fminf(color.x, 1.0f)));
    rgb8[3 * (size_t)p + 1] = (uint8_t)(255.99f * sqrt_exact(fminf(color.y, 1.0f)));
    rgb8[3 * (size_t)p + 2] = (uint8_t)(255.99f * sqrt_exact(fminf(color.z, 1.0f)));
    film.rngA[p] = make_uint4(xr.st.v0, xr.st.v1, xr.st.v2, xr.st.v3);
    film.rngB[p] = make_uint2(xr.st.v4, xr.st.d);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
static inline unsigned blocks_for(size_t n, int per_block) { return (unsigned)((n + (size_t)per_block - 1) / (size_t)per_block); }

uint32_t owned_pixel_count(int w, int h, int shard_mode, int rank, int world) {
    if (shard_mode != 1 || world <= 1) return (uint32_t)w * (uint32_t)h;
    uint32_t rows = 0;
    for (int band = rank; band * kBandRows < h; band += world) {
        int r0 = band * kBandRows, r1 = r0 + kBandRows;
        if (r1 > h) r1 = h;
        rows += (uint32_t)(r1 - r0);
    }
    return rows * (uint32_t)w;
}

void launch_rng_init(cudaStream_t st, const FilmView& film, unsigned long long seed_base, const uint32_t* jump_tables) {
    size_t npix = (size_t)film.width * film.height;
    if (!npix) return;
    k_rng_init<<<blocks_for(npix, kBlock), kBlock, 0, st>>>(film, seed_base, jump_tables);
}
void launch_slot_pixels(cudaStream_t st, uint32_t* slot_pixel, int w, int h, int shard_mode, int rank, int world, uint32_t n_owned) {
    if (!n_owned) return;
    k_slot_pixels<<<blocks_for(n_owned, kBlock), kBlock, 0, st>>>(slot_pixel, w, h, shard_mode, rank, world, n_owned);
}
void launch_raygen(cudaStream_t st, int rng_mode, const WaveState& ws, const FilmView& film, const CameraView& cam, const RenderParams& rp) {
    if (!rp.n_slots) return;
    unsigned g = blocks_for(rp.n_slots, kBlock);
    if (rng_mode == 0) k_raygen<0><<<g, kBlock, 0, st>>>(ws, film, cam, rp);
    else k_raygen<1><<<g, kBlock, 0, st>>>(ws, film, cam, rp);
}
void launch_extend(cudaStream_t st, const WaveState& ws, const SceneView& scene, int q, uint32_t qlen_bound, bool count) {
    if (!qlen_bound) return;
    unsigned g = blocks_for(qlen_bound, kBlock);
    if (count) k_extend<true><<<g, kBlock, 0, st>>>(ws, scene, q);
    else k_extend<false><<<g, kBlock, 0, st>>>(ws, scene, q);
}
void launch_shade(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film, const CameraView& cam,
                  const RenderParams& rp, int q, uint32_t qlen_bound) {
    if (!qlen_bound) return;
    unsigned g = blocks_for(qlen_bound, kBlock);
    if (rng_mode == 0) k_shade<0, false><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp, q);
    else k_shade<1, false><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp, q);
}
void launch_bounce(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film, const CameraView& cam,
                   const RenderParams& rp, int q, uint32_t qlen_bound) {
    if (!qlen_bound) return;
    unsigned g = blocks_for(qlen_bound, kBlock);
    if (rng_mode == 0) k_shade<0, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp, q);
    else k_shade<1, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp, q);
}
void launch_path_loop(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film, const CameraView& cam,
                      const RenderParams& rp, int q, uint32_t qlen_bound, int iters, int min_blocks) {
    if (!qlen_bound || iters <= 0) return;
    unsigned g = blocks_for(qlen_bound, kBlock);
    if (min_blocks == 4) {          // 64 registers, 4 CTAs / SM
        if (rng_mode == 0) k_path_loop<0, 4><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp, q, iters);
        else k_path_loop<1, 4><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp, q, iters);
    } else {                        // 80 registers, 3 CTAs / SM
        if (rng_mode == 0) k_path_loop<0, 3><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp, q, iters);
        else k_path_loop<1, 3><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp, q, iters);
    }
}
void launch_path_persist(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film, const CameraView& cam,
                         const RenderParams& rp, int sm_count, int min_blocks) {
    if (rp.spp <= 0 || !rp.n_owned) return;
    const unsigned long long items = rng_mode == 0 ? (unsigned long long)rp.n_owned : rp.total_paths;
    // resident CTAs: 3 x 256 threads at 80 registers (default), 4 x 256 at 64, or 7 x 128 at 72 (reserved[2] = 4 / 7);
    // reserved[2] = 1 / 2 launch only 1 / 2 of the three 256-thread CTAs per SM: fewer warps share an SM, so each pixel's
    // sequential sample chain advances faster (matters when a rank owns few pixels: DESIGN.md section 6)
    // bits 4-5 of reserved[2]: 1 = force the two-records-per-round triangle loop, 2 = force one record per round; 0 = by scene
    // size (worth it where the records come from L2 / HBM, not on cache-resident scenes: traverse.h, Traversal::step)
    const int pair_opt = (min_blocks >> 4) & 3;
    min_blocks &= 15;
    const bool pair = pair_opt == 1 || (pair_opt == 0 && scene.n_tris > 1024u);      // C3 at 41 K / 655 K / 10.5 M triangles: +2.3 / +2.7 / +6.1 %; C2 (32): -0.6 %
    int per_sm = min_blocks == 4 ? 4 : (min_blocks == 7 ? 7 : ((min_blocks == 1 || min_blocks == 2) ? min_blocks : 3));
    // default: with the sequential RNG and fewer than ~3 pixels per resident lane the frame is bounded by its longest pixel
    // chain, not by throughput -- two CTAs per SM finish such a frame sooner (C2, 1/8 of the film: 38 vs 43 ms)
    if (min_blocks == 0 && rng_mode == 0 && (unsigned long long)rp.n_owned < 3ull * (unsigned long long)sm_count * 3ull * kBlock) per_sm = 2;
    // counter RNG on a large scene: its build fits 64 registers without spills, and the latency-bound traversal gains from 32 warps
    // per SM (C3: 7 505 -> 8 046 M segments/s; the compat build spills 158 bytes at 64 registers and loses 7.5 %: profiles/r2y3_c3_minb4_pair.log)
    if (min_blocks == 0 && rng_mode != 0 && pair) per_sm = 4;
    const int block = min_blocks == 7 ? 128 : kBlock;
    unsigned g = (unsigned)std::min<unsigned long long>((unsigned long long)sm_count * per_sm, (items + block - 1) / block);
    if (scene.inst != nullptr) {
        // two-level scene: the cursor carries the object-space ray as well.  3 CTAs per SM at 80 registers (66 bytes of spills in the
        // compat build) beat 2 CTAs at 110: the kernel is latency-bound (ncu r2x on C5: issue slots 52 % busy at 16 warps per SM);
        // C5 on one GPU 1 163 -> 1 292 M segments/s, counter RNG 1 201 -> 1 422 (profiles/r2y_c5_ab.log).  ctas_per_sm = 2 keeps the old form.
        if (min_blocks == 0 || min_blocks == 4) {
            // default: 4 CTAs per SM at 64 registers (260 bytes of spills in the compat build, 74 with the counter RNG): C5 on one GPU
            // 1 293 -> 1 338 M segments/s, counter RNG 1 423 -> 1 440 (profiles/r2y2_c5_ab.log); ctas_per_sm = 3 / 2 keep the other forms.
            // (Keeping the cold per-pixel state -- running sum, pixel coordinates, counters: 8 registers -- in shared memory instead cut
            //  the spills to 196 bytes but gave 1 288: the 32 KB carve-out comes out of L1.  profiles/r2y4_c5_cold.log)
            g = (unsigned)std::min<unsigned long long>((unsigned long long)sm_count * 4, (items + kBlock - 1) / kBlock);
            if (rng_mode == 0) k_path_persist<0, kBlock, 4, false, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
            else k_path_persist<1, kBlock, 4, false, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
            return;
        }
        if (min_blocks != 2) {
            g = (unsigned)std::min<unsigned long long>((unsigned long long)sm_count * 3, (items + kBlock - 1) / kBlock);
            if (rng_mode == 0) k_path_persist<0, kBlock, 3, false, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
            else k_path_persist<1, kBlock, 3, false, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
            return;
        }
        g = (unsigned)std::min<unsigned long long>((unsigned long long)sm_count * 2, (items + kBlock - 1) / kBlock);
        if (rng_mode == 0) k_path_persist<0, kBlock, 2, false, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
        else k_path_persist<1, kBlock, 2, false, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
        return;
    }
    if (rng_mode == 0 && rp.shade.sampling_mode != 0 && scene.guide != nullptr) {
        // grid / MIS sampling (compat RNG only): the instantiation whose shading honours the per-primitive guide records
        g = (unsigned)std::min<unsigned long long>((unsigned long long)sm_count * std::min(per_sm, 3), (items + kBlock - 1) / kBlock);
        k_path_persist<0, kBlock, 3, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
        return;
    }
    if (per_sm == 4 && pair) {
        if (rng_mode == 0) k_path_persist<0, kBlock, 4, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
        else k_path_persist<1, kBlock, 4, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
    } else if (per_sm == 4) {
        if (rng_mode == 0) k_path_persist<0, kBlock, 4><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
        else k_path_persist<1, kBlock, 4><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
    } else if (per_sm == 7) {
        if (rng_mode == 0) k_path_persist<0, 128, 7><<<g, 128, 0, st>>>(ws, scene, film, cam, rp);
        else k_path_persist<1, 128, 7><<<g, 128, 0, st>>>(ws, scene, film, cam, rp);
    } else if (pair) {
        if (rng_mode == 0) k_path_persist<0, kBlock, 3, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
        else k_path_persist<1, kBlock, 3, false, true><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
    } else {
        if (rng_mode == 0) k_path_persist<0, kBlock, 3><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
        else k_path_persist<1, kBlock, 3><<<g, kBlock, 0, st>>>(ws, scene, film, cam, rp);
    }
}
void launch_cull_pixels(cudaStream_t st, FilmView film, const uint32_t* slot_base, uint32_t n_owned, CullRect rect, const uint32_t* row_prefix,
                        const uint32_t* skip, uint32_t weyl_add, uint32_t* slot_active) {
    if (!n_owned) return;
    k_cull_pixels<<<(n_owned + kBlock - 1) / kBlock, kBlock, 0, st>>>(film, slot_base, n_owned, rect, row_prefix, skip, weyl_add, slot_active);
}
void launch_path_flat(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FlatScene& fs, const FilmView& film,
                      const CameraView& cam, const RenderParams& rp, int sm_count, int per_sm) {
    if (rp.spp <= 0 || !rp.n_owned) return;
    const unsigned long long items = rng_mode == 0 ? (unsigned long long)rp.n_owned : rp.total_paths;
    if (per_sm != 1 && per_sm != 2 && per_sm != 4) per_sm = 3;      // 1: probe only (how fast does a lone sample chain run?)
    // few pixels per resident lane (a tile-sharded frame on many GPUs): fewer warps per SM, so every pixel's sequential
    // sample chain advances faster (see launch_path_persist)
    unsigned g = (unsigned)std::min<unsigned long long>((unsigned long long)sm_count * per_sm, (items + kBlock - 1) / kBlock);
    if (per_sm == 4) {
        if (rng_mode == 0) k_path_flat<0, kBlock, 4><<<g, kBlock, 0, st>>>(ws, scene, fs, film, cam, rp);
        else k_path_flat<1, kBlock, 4><<<g, kBlock, 0, st>>>(ws, scene, fs, film, cam, rp);
    } else {
        if (rng_mode == 0) k_path_flat<0, kBlock, 3><<<g, kBlock, 0, st>>>(ws, scene, fs, film, cam, rp);
        else k_path_flat<1, kBlock, 3><<<g, kBlock, 0, st>>>(ws, scene, fs, film, cam, rp);
    }
}
#if defined(PTX_EXPERIMENTAL)
template <int RNG, int BLOCK, int PPT, int MINB>
static void launch_cta_variant(cudaStream_t st, const WaveState& ws, const SceneView& scene, const FilmView& film, const CameraView& cam,
                               const RenderParams& rp, int sm_count, unsigned long long items) {
    constexpr int P = BLOCK * PPT;
    const size_t smem = sizeof(CtaPaths<P, 4 * P>);
    auto kern = k_path_cta<RNG, BLOCK, PPT, MINB>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    unsigned g = (unsigned)std::min<unsigned long long>((unsigned long long)sm_count * MINB, (items + P - 1) / P);
    kern<<<g, BLOCK, smem, st>>>(ws, scene, film, cam, rp);
}
// variant: 0 = 128 threads x 1 path, 8 CTAs / SM (the fastest measured); 1 = 256 threads x 2 paths, 3 CTAs / SM
void launch_path_cta(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film, const CameraView& cam,
                     const RenderParams& rp, int sm_count, int variant) {
    if (rp.spp <= 0 || !rp.n_owned) return;
    const unsigned long long items = rng_mode == 0 ? (unsigned long long)rp.n_owned : rp.total_paths;
#define PTX_CTA_LAUNCH(B, T, M)                                                                              \
    do {                                                                                                     \
        if (rng_mode == 0) launch_cta_variant<0, B, T, M>(st, ws, scene, film, cam, rp, sm_count, items);    \
        else launch_cta_variant<1, B, T, M>(st, ws, scene, film, cam, rp, sm_count, items);                  \
    } while (0)
    if (variant == 1) PTX_CTA_LAUNCH(256, 2, 3);
    else PTX_CTA_LAUNCH(128, 1, 8);
#undef PTX_CTA_LAUNCH
}
#endif  // PTX_EXPERIMENTAL
void launch_build_guide(cudaStream_t st, const float* rgb, size_t n_prims, float* out) {
    if (!n_prims) return;
    k_build_guide<<<blocks_for(n_prims, kBlock), kBlock, 0, st>>>(rgb, n_prims, out);
}
void launch_pixel_cost(cudaStream_t st, const SceneView& scene, const FlatScene* fs, const CameraView& cam, const ShadeParams& sp, const uint32_t* slot_base,
                       uint32_t n_owned, uint32_t* keys, uint32_t* vals) {
    if (!n_owned) return;
    if (fs) k_pixel_cost<true><<<blocks_for(n_owned, kBlock), kBlock, 0, st>>>(scene, *fs, cam, sp, slot_base, n_owned, keys, vals);
    else { FlatScene none{}; k_pixel_cost<false><<<blocks_for(n_owned, kBlock), kBlock, 0, st>>>(scene, none, cam, sp, slot_base, n_owned, keys, vals); }
}
void launch_queue_keys(cudaStream_t st, const WaveState& ws, const SceneView& scene, int q, uint32_t n_bound, bool by_material, const float lo[3],
                       const float hi[3], uint32_t* keys) {
    if (!n_bound) return;
    float3 l = make_float3(lo[0], lo[1], lo[2]);
    float3 ie = make_float3(32.0f / fmaxf(hi[0] - lo[0], 1e-20f), 32.0f / fmaxf(hi[1] - lo[1], 1e-20f), 32.0f / fmaxf(hi[2] - lo[2], 1e-20f));
    k_queue_keys<<<blocks_for(n_bound, kBlock), kBlock, 0, st>>>(ws, scene, q, n_bound, by_material ? 1 : 0, l, ie, keys);
}
void launch_compact(cudaStream_t st, const WaveState& ws, int q, uint32_t qlen_bound, uint32_t* scratch) {
    if (!qlen_bound) return;
    unsigned nb = blocks_for(qlen_bound, kCompactChunk);
    k_compact_count<<<nb, kBlock, 0, st>>>(ws, q, scratch);
    k_compact_scan<<<1, 1024, 0, st>>>(scratch, nb);
    k_compact_scatter<<<nb, kBlock, 0, st>>>(ws, q, scratch);
    k_compact_commit<<<1, 1, 0, st>>>(ws, scratch, nb);
}
void launch_resolve(cudaStream_t st, const FilmView& film, float inv_spp, uint8_t* rgb8, int flip_y) {
    size_t npix = (size_t)film.width * film.height;
    if (!npix) return;
    k_resolve<<<blocks_for(npix, kBlock), kBlock, 0, st>>>(film, inv_spp, rgb8, flip_y);
}
void launch_film_pack_owned(cudaStream_t st, const FilmView& film, const uint32_t* slot_pixel, uint32_t n_owned, float4* dst) {
    if (!n_owned) return;
    k_film_pack_owned<<<blocks_for(n_owned, kBlock), kBlock, 0, st>>>(film, slot_pixel, n_owned, dst);
}
void launch_film_unpack_all(cudaStream_t st, const FilmView& film, const float4* gathered, uint32_t per_rank, int world) {
    size_t npix = (size_t)film.width * film.height;
    if (!npix) return;
    k_film_unpack_all<<<blocks_for(npix, kBlock), kBlock, 0, st>>>(film, gathered, per_rank, world);
}
void launch_film_scale(cudaStream_t st, const FilmView& film, float inv_spp, float* linear_rgb) {
    size_t npix = (size_t)film.width * film.height;
    if (!npix) return;
    k_film_scale<<<blocks_for(npix, kBlock), kBlock, 0, st>>>(film, inv_spp, linear_rgb);
}
void launch_primary_ids(cudaStream_t st, const SceneView& scene, const CameraView& cam, int w, int h, int32_t* prim, float* t) {
    size_t n = (size_t)w * h;
    if (!n) return;
    if (scene.inst) k_primary_ids_inst<<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, cam, w, h, prim, t);
    else k_primary_ids<<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, cam, w, h, prim, t);
}
void launch_trace_rays(cudaStream_t st, const SceneView& scene, const float* rays, size_t n, float t_min, int32_t* prim, float* t) {
    if (!n) return;
    if (scene.inst) k_trace_rays_inst<<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, rays, n, t_min, prim, t);
    else k_trace_rays<<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, rays, n, t_min, prim, t);
}
void launch_offset_nodes(cudaStream_t st, u32x4* nodes, uint32_t n_nodes, uint32_t node_off, uint32_t tri_off) {
    if (!n_nodes) return;
    k_offset_nodes<<<blocks_for(n_nodes, kBlock), kBlock, 0, st>>>(nodes, n_nodes, node_off, tri_off);
}

// ---------------------------------------------------------------------------------------------
// Allocation-free closest-hit query (ptx_trace_rays_noalloc; no counterpart in the reference -- added
// for callers that must not call cudaMalloc: code under CUDA graph capture, tight serving loops).
// ptx_trace_rays stages rays / ids / distances in device buffers it cudaMallocs per call; this variant runs the same
// k_trace_rays kernel on a file-scope device array instead.  Every GPU that loads the library holds the array (2 MB) until
// exit.  All calls share it: calls that overlap on the GPU (two contexts of one device, two host threads) RACE on it.
// Nothing in it has to start at zero: a call writes rays [0, n) before the kernel reads them and the kernel writes ids / t
// [0, n) before they are copied out, so the previous call's contents beyond n are never read.
// g_noalloc_scratch is a HOST-SIDE STAND-IN in host code, not a device address: the address the kernel gets comes from
// cudaGetSymbolAddress, resolved once per device with that device current.
// ---------------------------------------------------------------------------------------------
constexpr size_t kNoAllocBytesPerRay = 6 * sizeof(float) + sizeof(int32_t) + sizeof(float);      // ray + prim id + t
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocMaxRays * kNoAllocBytesPerRay];

namespace {
constexpr int kNoAllocMaxDevices = 64;
void* g_noalloc_ptr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mutex;
}  // namespace

cudaError_t noalloc_scratch_address(int device, void** out) {
    *out = nullptr;
    if (device < 0 || device >= kNoAllocMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;          // the caller makes the device current first (api.cu does)
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

cudaError_t trace_rays_noalloc(cudaStream_t st, int device, const SceneView& scene, const float* rays_host, size_t n, float t_min,
                               int32_t* prim_host, float* t_host) {
    if (n > kNoAllocMaxRays) return cudaErrorInvalidValue;
    void* base = nullptr;
    cudaError_t e = noalloc_scratch_address(device, &base);
    if (e != cudaSuccess) return e;
    float* d_rays = reinterpret_cast<float*>(base);                                                   // [kNoAllocMaxRays][6]
    int32_t* d_ids = reinterpret_cast<int32_t*>(static_cast<unsigned char*>(base) + kNoAllocMaxRays * 6 * sizeof(float));
    float* d_t = reinterpret_cast<float*>(static_cast<unsigned char*>(base) + kNoAllocMaxRays * 7 * sizeof(float));
    e = cudaMemcpyAsync(d_rays, rays_host, n * 6 * sizeof(float), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    if (scene.inst) k_trace_rays_inst<<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, d_rays, n, t_min, d_ids, d_t);
    else k_trace_rays<<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, d_rays, n, t_min, d_ids, d_t);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(prim_host, d_ids, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t_host, d_t, n * sizeof(float), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    return e;
}

void launch_occluded(cudaStream_t st, const SceneView& scene, const float* segs, size_t n, uint8_t* out) {
    if (!n) return;
    if (scene.inst) k_occluded_inst<<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, segs, n, out);
    else k_occluded<<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, segs, n, out);
}
void launch_visibility(cudaStream_t st, const SceneView& scene, const FilmView& film, const CameraView& cam, int spp, const float* radiosity, uint8_t* rgb8) {
    size_t n = (size_t)cam.width * cam.height;
    if (!n || spp <= 0) return;
    if (scene.inst) k_visibility<true><<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, film, cam, spp, radiosity, rgb8);
    else k_visibility<false><<<blocks_for(n, kBlock), kBlock, 0, st>>>(scene, film, cam, spp, radiosity, rgb8);
}

}  // namespace ptx
