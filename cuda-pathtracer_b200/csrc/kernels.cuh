// kernels.cuh -- launch interface of the sm_100a wavefront kernels (kernels.cu).
//
// Pipeline (replaces the one-thread-per-pixel megakernel `render`, rendering/integrator.h:371-408):
//
//   rng_init     once per film      XORWOW state per pixel == curand_init(seed+pix, pix, 0)   (render_init, :274-280)
//   raygen       once per frame     first camera ray of every slot, queue = identity
//   extend       per iteration      closest hit of every live slot's ray through the BVH8      (Scene::intersect)
//   shade        per iteration      emission -> film, Russian roulette, albedo, cosine sample; a path that
//                                   ends is REGENERATED in place with the pixel's next sample (so warps stay full);
//                                   a slot whose pixel has no samples left is marked dead
//   compact      when worthwhile    stable, warp-aggregated compaction of the live-slot queue
//   resolve      once per frame     mean, Reinhard, gamma, RGB8                                 (:393-407)
//
// Path state is SoA, three 16-byte records per slot so every access is one coalesced 128-bit
// load/store per lane:
//   A = { o.x, o.y, o.z, d.x }   B = { d.y, d.z, depth|flags, beta.r }   C = { beta.g, beta.b, hit_t, hit_prim }
#pragma once
#include "cull.h"
#include <cuda_runtime.h>

#include "flat.h"
#include "layout.h"
#include "shade.h"
#include "traverse_inst.h"

namespace ptx {

constexpr uint32_t kDeadFlag = 0x80000000u;

struct WaveState {
    float4* A;
    float4* B;
    float4* C;
    float4* E;                // path-loop pipeline: { L.r, L.g, L.b, - } radiance of the path in flight
    uint2* D;                 // philox mode: { pixel, sample index }; compat mode: { pixel, samples done }
    uint32_t* queue[2];       // live slot indices (ping-pong for compaction)
    uint32_t* slot_pixel;     // slot -> pixel for the pixels this rank owns
    uint32_t* progress;       // persistent kernel, compat RNG: sample chunks finished, per pixel of the film (zeroed per frame)
    // device counters (one 64-byte block): see CounterBlock
    unsigned long long* counters;
    uint32_t capacity;        // slots allocated
};

// indices into WaveState::counters
enum { CTR_SEGMENTS = 0, CTR_PATHS = 1, CTR_LIVE = 2, CTR_QLEN = 3, CTR_NEXT_PATH = 4, CTR_NODE_VISITS = 5, CTR_TRI_TESTS = 6, CTR_SLOT_LAUNCHES = 7, CTR_COUNT = 8 };

struct FilmView {
    float4* accum;            // per pixel: sum r, g, b, (unused)
    uint4* rngA;              // per pixel: v0..v3
    uint2* rngB;              // per pixel: v4, d
    int width, height;
};

constexpr int kMaxChunks = 40;
struct RenderParams {
    ShadeParams shade;
    int spp;
    uint32_t n_slots;         // owned pixels (compat) / pool size (philox)
    uint32_t n_owned;         // pixels owned by this rank
    unsigned long long total_paths;   // philox: n_owned * spp
    uint32_t sample_begin;    // philox sample-range sharding: first global sample index
    int shade_batch;          // persistent kernel: waiting lanes per warp that trigger a shading round (1 = shade at once)
    int n_chunks;             // persistent kernel, compat RNG: work items per pixel (consecutive sample ranges, see k_path_persist)
    uint32_t chunk_start[kMaxChunks + 1];   // chunk c covers samples [chunk_start[c], chunk_start[c + 1]); chunk_start[n_chunks] = spp
    unsigned long long seed;
};

void launch_rng_init(cudaStream_t st, const FilmView& film, unsigned long long seed_base, const uint32_t* jump_tables);
void launch_raygen(cudaStream_t st, int rng_mode, const WaveState& ws, const FilmView& film, const CameraView& cam,
                   const RenderParams& rp);
void launch_extend(cudaStream_t st, const WaveState& ws, const SceneView& scene, int q, uint32_t qlen_bound, bool count);
void launch_shade(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film,
                  const CameraView& cam, const RenderParams& rp, int q, uint32_t qlen_bound);
void launch_bounce(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film,
                   const CameraView& cam, const RenderParams& rp, int q, uint32_t qlen_bound);
// path-loop pipeline: every live slot runs up to `iters` bounces (extend + shade + regeneration) with its
// state in registers; state goes through HBM once per launch instead of once per bounce
void launch_path_loop(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film,
                      const CameraView& cam, const RenderParams& rp, int q, uint32_t qlen_bound, int iters, int min_blocks);
void launch_path_persist(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film, const CameraView& cam,
                         const RenderParams& rp, int sm_count, int min_blocks);
// tiny scenes (pipeline 6): the persistent loop with flat_closest() instead of the BVH8 traversal (flat.h); per_sm = resident CTAs per SM (2, 3, 4)
void launch_path_flat(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FlatScene& fs, const FilmView& film,
                      const CameraView& cam, const RenderParams& rp, int sm_count, int per_sm);
// CTA wavefront (pipeline 5): one launch per frame, paths in shared memory, four dense phases per iteration (kernels.cu)
#if defined(PTX_EXPERIMENTAL)
void launch_path_cta(cudaStream_t st, int rng_mode, const WaveState& ws, const SceneView& scene, const FilmView& film, const CameraView& cam,
                     const RenderParams& rp, int sm_count, int variant);
#endif
// guide records (kGuideFloats floats per primitive, shade.h) from n_prims x 256 x 3 device floats of radiosity
void launch_build_guide(cudaStream_t st, const float* rgb, size_t n_prims, float* out);
// cost probe for the work order of the persistent kernels (kernels.cu k_pixel_cost): keys[i] = 255 - segments of 2 probe paths of
// pixel slot_base[i], vals[i] = that pixel; fs = the flat unit list or nullptr (BVH8)
void launch_pixel_cost(cudaStream_t st, const SceneView& scene, const FlatScene* fs, const CameraView& cam, const ShadeParams& sp, const uint32_t* slot_base,
                       uint32_t n_owned, uint32_t* keys, uint32_t* vals);
// sort keys of queue[q][0, n_bound): ray keys (origin cell + direction octant) or material keys (after extend); kernels.cu
void launch_queue_keys(cudaStream_t st, const WaveState& ws, const SceneView& scene, int q, uint32_t n_bound, bool by_material, const float lo[3],
                       const float hi[3], uint32_t* keys);
// compacts queue[q] -> queue[q^1]; scratch holds per-block counts (>= ceil(qlen_bound/1024)+1 words)
void launch_compact(cudaStream_t st, const WaveState& ws, int q, uint32_t qlen_bound, uint32_t* scratch);
void launch_resolve(cudaStream_t st, const FilmView& film, float inv_spp, uint8_t* rgb8, int flip_y);
// tile-sharded film combine: pack the owned pixels' sums (slot order) / unpack all ranks' slices (per_rank float4 each) into the film
void launch_film_pack_owned(cudaStream_t st, const FilmView& film, const uint32_t* slot_pixel, uint32_t n_owned, float4* dst);
void launch_film_unpack_all(cudaStream_t st, const FilmView& film, const float4* gathered, uint32_t per_rank, int world);
void launch_film_scale(cudaStream_t st, const FilmView& film, float inv_spp, float* linear_rgb);
void launch_primary_ids(cudaStream_t st, const SceneView& scene, const CameraView& cam, int w, int h, int32_t* prim, float* t);
void launch_trace_rays(cudaStream_t st, const SceneView& scene, const float* rays, size_t n, float t_min, int32_t* prim, float* t);
// allocation-free variant of the closest-hit query: staging in a file-scope device array of kernels.cu (see there); copies the
// rays in, runs k_trace_rays, copies ids / distances out, synchronises the stream.  `device` must be the current device.
constexpr size_t kNoAllocMaxRays = 65536;
cudaError_t trace_rays_noalloc(cudaStream_t st, int device, const SceneView& scene, const float* rays_host, size_t n, float t_min,
                               int32_t* prim_host, float* t_host);
void launch_offset_nodes(cudaStream_t st, u32x4* nodes, uint32_t n_nodes, uint32_t node_off, uint32_t tri_off);   // two-level scene assembly
void launch_occluded(cudaStream_t st, const SceneView& scene, const float* segs, size_t n, uint8_t* out);
void launch_visibility(cudaStream_t st, const SceneView& scene, const FilmView& film, const CameraView& cam, int spp, const float* radiosity,
                       uint8_t* rgb8);
void launch_slot_pixels(cudaStream_t st, uint32_t* slot_pixel, int w, int h, int shard_mode, int rank, int world, uint32_t n_owned);
// Empty-space culling of whole pixels (api.cu, ptx_render): the owned pixels inside the screen rectangle [x0, x1) x [y0, y1) are
// compacted, in slot order, into slot_active (position = row_prefix[y] * (x1 - x0) + x - x0, row_prefix = owned rows in [y0, y));
// every pixel outside has its XORWOW stream advanced by the draws its samples would have consumed (skip = T^n as 800 words, the
// Weyl counter by weyl_add), or nothing when skip is NULL (counter RNG).
void launch_cull_pixels(cudaStream_t st, FilmView film, const uint32_t* slot_base, uint32_t n_owned, CullRect rect, const uint32_t* row_prefix,
                        const uint32_t* skip, uint32_t weyl_add, uint32_t* slot_active);

// pixels owned by `rank` under tile sharding: rows are dealt in bands of kBandRows
constexpr int kBandRows = 4;
uint32_t owned_pixel_count(int w, int h, int shard_mode, int rank, int world);

}  // namespace ptx
