/* include/ptx.h -- C ABI of the B200-native path-tracing core ("ptx" = path-tracer core, the
 * prefix proposed in SURVEY.md section 8(b); nothing to do with NVIDIA's PTX ISA).
 *
 * Drop-in boundary for the render loop of USharma002/CUDA-PathTracer.  The reference has no FFI
 * layer; its "API" is the set of C++ structs and free functions that SceneState::loadScene,
 * RenderState::allocateBuffers and renderFrame call directly.  Each entry point below names the
 * reference call (file:line under /root/reference) it replaces.  Conventions:
 *   - plain C, opaque handles, plain pointers and sizes; no C++/torch types cross this boundary;
 *   - every function returns 0 on success or a negative ptx_status; nothing throws;
 *     ptx_last_error() gives the text (the reference prints to stderr and carries on,
 *     application_state.h:394-397; cudaMallocSafe throws, utils/cuda_utils.h:53-60);
 *   - a ptx_ctx is one CUDA device + one stream; calls on one ctx are not thread-safe;
 *   - all compute runs in hand-written sm_100a CUDA kernels; there is NO CPU fallback: without a
 *     usable CUDA device ptx_ctx_create fails with PTX_ERR_NO_DEVICE.
 */
#ifndef PTX_H
#define PTX_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PTX_OK = 0,
    PTX_ERR_INVALID = -1,     /* bad argument */
    PTX_ERR_IO = -2,          /* file missing / unreadable / no primitives (loadOBJ returns false) */
    PTX_ERR_CUDA = -3,        /* CUDA runtime error; text in ptx_last_error */
    PTX_ERR_NO_DEVICE = -4,   /* no CUDA device: the product has no CPU path */
    PTX_ERR_UNSUPPORTED = -5, /* a combination of options that is not implemented (e.g. grid sampling with the counter RNG) */
    PTX_ERR_NOMEM = -6
} ptx_status;

typedef struct ptx_ctx ptx_ctx;
typedef struct ptx_scene ptx_scene;
typedef struct ptx_film ptx_film;

/* ---- context ------------------------------------------------------------------------------- */
int ptx_ctx_create(int device, ptx_ctx** out);
void ptx_ctx_destroy(ptx_ctx* ctx);
const char* ptx_last_error(const ptx_ctx* ctx);           /* ctx may be NULL: last error of this thread */
int ptx_sync(ptx_ctx* ctx);                                /* cudaDeviceSynchronize, application.h:199 */
int ptx_ctx_device_info(const ptx_ctx* ctx, int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem);
/* benchmarking aid: overwrite a 256 MB scratch buffer on the context's stream so that nothing of the previous frame is
 * left in the 126 MB L2 (the persistent kernel's per-frame inputs, film + RNG state, are smaller than the L2) */
int ptx_ctx_flush_l2(ptx_ctx* ctx);
const char* ptx_version(void);
/* CUDA-event stopwatch on the ctx stream (what bench.py times steps with) and the raw cudaStream_t */
int ptx_timer_start(ptx_ctx* ctx);
int ptx_timer_stop(ptx_ctx* ctx, float* ms);               /* records, synchronises, returns elapsed ms */
int ptx_ctx_stream(const ptx_ctx* ctx, void** stream);

/* ---- scene --------------------------------------------------------------------------------- */
/* One reference `Primitive` minus its radiosity payload (rendering/primitive.h:21-29,
 * triangle.h:98-105, quad.h:134-140).  type 0: v[0..2] = v0 v1 v2; type 1: v00 v10 v11 v01. */
typedef struct {
    int32_t type;
    float v[4][3];
    float normal[3];
    float bsdf[3];
    float Le[3];
    int32_t material_type;    /* 0 = Lambertian (all the reference has); 1 = perfect mirror (extension) */
} ptx_primitive;

typedef struct {
    float bsdf[3];            /* Kd (or Kr for mirrors) */
    float Le[3];              /* Ke */
    int32_t type;             /* 0 diffuse, 1 mirror */
    int32_t reserved;
} ptx_material;

typedef struct {
    int32_t convert_quads;    /* AppConfig::convert_quads_to_triangles (application_state.h:271) */
    int32_t subdivision;      /* AppConfig::subdivision_step: 1->4 splits per level (form_factors.h:520-574) */
    int64_t triangle_cap;     /* .pbrt only: 0 = no cap (the reference substitutes a box above 2 M, pbrt_loader.h:205) */
    int32_t mirror_from_mtl;  /* extension: MTL `illum 3|5` => mirror with Kr = Ks.  0 keeps reference semantics */
    int32_t host_preprocess;  /* 1 = run convert_quads / subdivision on the host (cross-check); default 0 = CUDA kernels
                                 feeding the device BVH builder (ptx_scene_load only; ptx_parse_scene_file is host-only) */
    int32_t keep_instances;   /* .pbrt, ptx_scene_load only: 1 = keep ObjectInstance two-level (every object stored once in object
                                 space, a TLAS over the instances; the ray is transformed at the instance) instead of flattening
                                 like the reference (pbrt_loader.h:197, 282-344).  Same primitive numbering; hit distances agree
                                 with the flattened scene to rounding, not bit for bit.  Ignored (flattened) for files without
                                 instances or with nested instancing, and with convert_quads / subdivision / triangle_cap.
                                 A two-level scene renders with BSDF sampling (pipeline 0 / 3) and answers ptx_trace_primary_ids,
                                 ptx_trace_rays(_noalloc), ptx_trace_occluded and ptx_render_visibility; grid / MIS sampling and
                                 ptx_scene_build_accel are not available for it. */
    int32_t reserved;
} ptx_load_opts;

typedef struct {
    int32_t builder;          /* 0 = GPU binned-SAH (default); 1 = host binned-SAH (cross-check) */
    int32_t sah_bins;         /* 0 = default (16) */
    int32_t max_leaf;         /* triangles per BVH8 leaf slot, 1..3; 0 = default (3) */
    int32_t reserved;
} ptx_build_opts;

typedef struct {
    int64_t n_primitives;     /* reference numbering: one per OBJ face / PBRT triangle */
    int64_t n_triangles;      /* quads contribute two */
    int64_t n_nodes;          /* BVH8 nodes (80 B each) */
    int64_t n_materials;
    int64_t node_bytes, triangle_bytes, shading_bytes;
    float bounds_min[3], bounds_max[3];
    float sah_cost;           /* SAH cost of the BVH8, root area normalised */
    float build_ms;           /* device time of the last build */
    int32_t builder;
    int32_t max_depth;
    float preprocess_ms;      /* device time of convert-quads / subdivision / packing when they ran on the GPU, else 0 */
    int32_t preprocess_on_device;
    int32_t n_instances;      /* > 0: two-level scene (keep_instances): instances in the TLAS; n_triangles is then the FLATTENED count */
    int32_t reserved2;
    int64_t stored_triangles; /* triangle records actually stored (two-level: every object once) */
    float build_stage_ms[4];  /* device builder: upload + boxes + Morton sort | binned-SAH levels | leaf order | BVH8 collapse */
    int32_t flat_units;       /* > 0: the scene is tiny enough for the flat unit list (<= 20 planar quads / triangle pairs):
                                 ptx_render's default pipeline then runs without a tree (pipeline 6) */
    int32_t reserved;
} ptx_scene_stats;

/* replaces loadScene/loadOBJ/loadPBRT + RayTracingManager::buildAccelStructure + Scene(...)
 * (utils/file_manager.h:93, 291-315; utils/pbrt_loader.h:178; ray_tracing_backend.h:81-129;
 * scene.h:23-27; driver application_state.h:367-464).  Builds the acceleration structure. */
int ptx_scene_load(ptx_ctx* ctx, const char* path, const ptx_load_opts* opts, ptx_scene** out);
/* host-side parse only (no device work): what loadOBJ returns, in the reference's numbering */
int ptx_parse_scene_file(const char* path, const ptx_load_opts* opts, ptx_primitive** prims_out, int64_t* n_out);
void ptx_free_primitives(ptx_primitive* prims);
/* scene from an in-memory primitive list (what a caller holding the reference's Primitive[] passes) */
int ptx_scene_from_primitives(ptx_ctx* ctx, const ptx_primitive* prims, int64_t n, const ptx_build_opts* bopts,
                              ptx_scene** out);
/* packed triangle soup: v = 9 floats per triangle; normals NULL => geometric; mat_id indexes mats */
int ptx_scene_from_triangles(ptx_ctx* ctx, const float* v, const float* normals, const uint32_t* mat_id,
                             const ptx_material* mats, size_t n_tri, size_t n_mat, const ptx_build_opts* bopts,
                             ptx_scene** out);
int ptx_scene_build_accel(ptx_scene* scene, const ptx_build_opts* opts);     /* rebuild with other options */
int ptx_scene_stats_get(const ptx_scene* scene, ptx_scene_stats* out);
int ptx_scene_get_primitives(const ptx_scene* scene, ptx_primitive* out, int64_t capacity);   /* returns the count (>= 0) */
void ptx_scene_destroy(ptx_scene* scene);
/* Grid-guided sampling data: SceneState::precomputeCDFs (application_state.h:492-585).  rgb_host = per reference primitive
 * (ptx_scene_get_primitives numbering) the 16 x 16 RGB grid the radiosity solver leaves in Triangle/Quad::radiosity_grid
 * (triangle.h:108, quad.h:143; row = theta band, rows 0..7 = upper hemisphere, column = phi band), n_primitives * 256 * 3
 * floats.  A CUDA kernel turns them into luminance pdf + marginal / conditional CDFs (Grid::initFromRadiosity + buildCDFs,
 * grid.h:62-137) and ptx_render's grid / MIS sampling modes read them.  The solver itself is outside this library.
 * NULL removes the data.  mis_bsdf_fraction = Scene::mis_bsdf_fraction (scene.h:215; the reference's default is 0.5). */
int ptx_scene_set_radiosity_grids(ptx_scene* scene, const float* rgb_host, float mis_bsdf_fraction);
/* the records built from them, 272 floats per primitive: pdf[8][16], row_cdfs[8][16], marginal_cdf[8], total_weight,
 * is_valid (as uint32), 6 x padding (parity tests) */
int ptx_scene_get_guide_records(const ptx_scene* scene, int64_t first, int64_t n, float* out_host);

/* ---- camera: the reference's Sensor (rendering/sensor.h:14-101), same field names ------------ */
typedef struct {
    float radius, yaw, pitch;
    float vup[3], lookat[3];
    float origin[3], lower_left_corner[3], horizontal[3], vertical[3];
    float vfov, aspect;
    int32_t image_width, image_height;
} ptx_camera;

int ptx_camera_set(ptx_camera* cam, const float eye[3], const float at[3], const float up[3], float vfov_deg,
                   float aspect);                                           /* Sensor ctor, sensor.h:16-29 */
int ptx_camera_update(ptx_camera* cam);                                     /* updateCamera, sensor.h:38-51 */
int ptx_camera_orbit(ptx_camera* cam);                                      /* updateCameraOrbit, sensor.h:56-67 */
/* RenderState default camera + allocateBuffers aspect + renderFrame's orbit update
 * (application_state.h:88, 106-109; application.h:161-162) */
int ptx_camera_default(ptx_camera* cam, int width, int height);

/* ---- film: RenderState::allocateBuffers (application_state.h:91-123) ------------------------- */
int ptx_film_create(ptx_ctx* ctx, int width, int height, ptx_film** out);   /* also runs render_init */
int ptx_film_reset(ptx_film* film, uint64_t seed_base);                     /* re-seed (reference: 2023) + clear */
int ptx_film_clear(ptx_film* film);                                         /* clear accumulation only */
/* Reinhard + gamma 1/2.2 + 255.99 quantisation of integrator.h:393-407; row 0 = bottom unless flip_y.
 * dst_is_device with dst == NULL resolves into the film's own device buffer (result stays in HBM). */
int ptx_film_resolve_rgb8(ptx_film* film, uint8_t* dst, int dst_is_device, int flip_y);
int ptx_film_read_linear(ptx_film* film, float* rgb_host);                  /* w*h*3 pre-tonemap means */
int ptx_film_write(ptx_film* film, const char* path);                        /* by extension: .png / .ppm (RGB8, top row first) or
                                                                                .pfm (linear radiance); replaces "Save PNG", ui_windows.h:195-210 */
/* host-only PNG writer for an RGB8 image the caller already holds (e.g. the reference's h_image, flip_y = 1 like
 * stbi_flip_vertically_on_write(1) + stbi_write_png, ui_windows.h:205-208).  Needs no GPU. */
int ptx_write_png_rgb8(const char* path, int width, int height, const uint8_t* rgb_host, int flip_y);
/* raw accumulation buffer (float4 per pixel: sum r,g,b + sample count) for NCCL film combine */
/* Checkpoint / resume (no counterpart in the reference: its image and curandState array live in device memory only).
 * ptx_film_save writes the running sums, the per-pixel XORWOW streams and the sample counters; ptx_film_load restores them into a film
 * of the same size, after which ptx_render(accumulate = 1) continues the progressive render exactly where it stopped (same film,
 * same streams, bit for bit).  PTX_ERR_IO (missing, not a checkpoint, truncated) / PTX_ERR_INVALID (size mismatch) on failure; the film is untouched then. */
int ptx_film_save(ptx_film* film, const char* path);
int ptx_film_load(ptx_film* film, const char* path);
int ptx_film_device_buffer(ptx_film* film, void** dev_ptr, size_t* bytes);
int ptx_film_set_sample_count(ptx_film* film, float spp_total);             /* after an external reduce */
void ptx_film_destroy(ptx_film* film);

/* ---- render: renderFrame (application.h:157-216) -> render<<<>>> (integrator.h:371-408) ------- */
enum { PTX_RNG_XORWOW_COMPAT = 0, PTX_RNG_PHILOX = 1 };
/* SamplingMode (rendering/render_config.h:41-47).  integrator.h:226-262: BSDF = cosine lobe; MIS = one-sample MIS between
 * the hit primitive's grid and the cosine lobe (sampleMIS, integrator.h:112-166); the other three = the grid alone.  Without
 * grids (ptx_scene_set_radiosity_grids) or on a primitive whose grid is empty every mode falls back to the cosine lobe. */
enum { PTX_SAMPLING_BSDF = 0, PTX_SAMPLING_FORMFACTOR = 1, PTX_SAMPLING_RADIOSITY = 2, PTX_SAMPLING_MIS = 3, PTX_SAMPLING_TOPK = 4 };
enum { PTX_SHARD_NONE = 0, PTX_SHARD_TILES = 1, PTX_SHARD_SAMPLES = 2 };

typedef struct {
    int32_t spp;              /* AppConfig::spp */
    int32_t max_depth;        /* reference hard-codes 5 (integrator.h:389); 0 => 5 */
    int32_t rng_mode;         /* PTX_RNG_* ; compat reproduces curand_init(2023+pix, pix, 0) streams */
    int32_t sampling_mode;    /* PTX_SAMPLING_*; the grid / MIS modes need rng_mode = PTX_RNG_XORWOW_COMPAT and pipeline 0 / 3 */
    int32_t accumulate;       /* 0: each call is an independent frame like the reference; 1: progressive */
    int32_t shard_mode;       /* PTX_SHARD_* */
    int32_t shard_rank, shard_world;
    int32_t rr_start_depth;   /* 0 => 3  (depth > 2, integrator.h:207) */
    float rr_cap;             /* 0 => 0.95 */
    int32_t pipeline;         /* 0 = default: 6 when the scene is tiny enough to carry a flat unit list (ptx_scene_stats.flat_units > 0,
                                     BSDF sampling), else 3.
                                 3 = persistent path loop over the BVH8: one launch per frame; lanes pull pixels / paths from a counter,
                                     no path state in memory; with the compat RNG a pixel's samples are split into work items
                                     handed from lane to lane through the film (see `chunks`).
                                 6 = the same loop with the tree replaced by one straight-line pass over <= 20 planar units
                                     (csrc/flat.h); identical hits.
                                 Cross-checks kept for the parity tests (bit-identical images): 4 = wavefront (raygen / extend / shade /
                                 compact kernels), 1 = fused bounce kernel, 2 = path loop with a slot queue.  5 = CTA wavefront: an experiment
                                 that measured slower, compiled only with -DPTX_EXPERIMENTAL (PTX_ERR_UNSUPPORTED otherwise). */
    int32_t flags;            /* PTX_RENDER_* bits below */
    int32_t loop_iters;       /* pipeline 2 only: bounces per launch (0 = default) */
    int32_t ctas_per_sm;      /* pipelines 2 / 3 / 6: resident 256-thread CTAs per SM, 1..4 (0 = chosen from pixels per lane and spp);
                                 pipeline 3 also: 7 = 7 x 128 threads, + 16 / + 32 force two / one triangle records per loop round */
    int32_t chunks;           /* compat RNG, pipelines 3 / 5 / 6: work items per pixel.  0 = default (1/2, 1/4, ... 1/32, 1/32 of spp, or
                                 eight equal parts with the last one halved down to 1/32 when a rank owns few pixels per resident lane
                                 and spp >= 512); n > 0: n halving chunks; n < 0: -n equal parts.  Images do not depend on it. */
    int32_t shade_batch;      /* pipeline 3: waiting lanes per warp that trigger a shading round (0 = default 12) */
} ptx_render_opts;

enum {
    PTX_RENDER_STAGE_TIMING = 1,   /* fill ptx_render_stats.extend_ms / shade_ms (pipelines 1, 2, 4) */
    PTX_RENDER_COUNT_WORK = 2,     /* pipeline 4: instrumented extend kernel, fills node_visits / tri_tests */
    PTX_RENDER_SORT_RAYS = 8,      /* pipeline 4: sort the queue by (origin cell, direction octant) before every extend (ray reordering) */
    PTX_RENDER_SORT_MATERIAL = 16, /* pipeline 4: sort the queue by the material of the hit before every shade (material-sorted shading).
                                      Neither changes an image (a slot's path, stream and pixel are its own); measured in DESIGN.md */
    PTX_RENDER_NO_COST_ORDER = 32, /* pipelines 3 / 6, compat RNG: keep the work items in pixel order instead of handing out the
                                      longest-expected pixels first (a 2-path probe per pixel + a radix sort per frame); images are the
                                      same either way, only the frame's tail differs (A/B measurements) */
    PTX_RENDER_NO_BOUNDS_CULL = 64, /* pipelines 3 / 6: trace every pixel.  By default the pixels none of whose camera rays can reach the
                                      scene's bounding box (outside the screen rectangle of its projection, 2 pixels of margin) are not
                                      traced: every sample of such a pixel draws its two jitter numbers, misses and ends, so the pixel's
                                      XORWOW stream is advanced by 2 * spp draws (one GF(2) matrix product) and its film sum left alone --
                                      film and RNG state equal those of the traced frame bit for bit.  The culled paths are counted in
                                      ptx_render_stats.paths / segments and reported in culled_paths */
    PTX_RENDER_ASYNC = 4           /* pipelines 3 / 6: ptx_render returns without waiting for the GPU (nothing between the render
                                      and a following ptx_film_combine / resolve synchronises the host); the statistics are
                                      completed by the next ptx_render_stats_get, ptx_sync or ptx_render on this context */
};

typedef struct {
    uint64_t segments;        /* Scene::intersect-equivalent calls (path segments) of the last render */
    uint64_t paths;
    uint64_t iterations;      /* wavefront iterations */
    uint64_t kernel_launches; /* kernels launched by the last ptx_render */
    uint64_t compactions;
    float render_ms;          /* CUDA-event time of the last ptx_render on the ctx stream */
    float extend_ms, shade_ms, other_ms;   /* filled with PTX_RENDER_STAGE_TIMING */
    uint64_t node_visits, tri_tests;       /* filled with PTX_RENDER_COUNT_WORK */
    uint64_t slot_launches;                /* path loop: sum over launches of the live slots that entered the kernel */
    uint64_t culled_paths;                 /* of `paths` (and of `segments`: one each): resolved by whole-pixel culling, not traced */
} ptx_render_stats;

/* Host-side sharding arithmetic (no device work; what a multi-process driver and the CPU tests use to check that the
 * ranks' shares are disjoint and complete).  Tiles: bands of 4 image rows dealt round-robin; samples: contiguous ranges. */
int64_t ptx_shard_pixel_count(int width, int height, int shard_mode, int rank, int world);
int ptx_shard_owned_pixels(int width, int height, int shard_mode, int rank, int world, uint32_t* pixels_out, int64_t capacity);
int ptx_shard_sample_range(int spp, int shard_mode, int rank, int world, int32_t* first, int32_t* count);

int ptx_render(ptx_ctx* ctx, const ptx_scene* scene, const ptx_camera* cam, ptx_film* film,
               const ptx_render_opts* opts);
int ptx_render_stats_get(const ptx_ctx* ctx, ptx_render_stats* out);

/* ---- multi-GPU film combine (no counterpart in the reference, which is single-GPU: SURVEY.md 2.3) -------------------
 * One process (or thread) per GPU, each with its own ptx_ctx, the scene replicated; every rank renders its share of the frame
 * (ptx_render_opts.shard_mode / shard_rank / shard_world) and ptx_film_combine exchanges the films over NCCL (NVLink / NVSwitch)
 * on the context's stream.  Afterwards EVERY rank's film holds the whole frame (resolve / read / write as usual).
 *   PTX_SHARD_TILES   (compat RNG): a rank packs the pixels it owns, ONE ncclAllGather moves 1/G of the film per rank, a
 *                     kernel unpacks all slices; the result equals the single-GPU film bit for bit.
 *   PTX_SHARD_SAMPLES (counter RNG): in-place ncclAllReduce(sum) of the running sums; the sample count becomes the total.
 * No host synchronisation happens in ptx_film_combine; with PTX_RENDER_ASYNC none happens between render and combine either.
 * NCCL is loaded at run time (dlopen of $PTX_NCCL_LIB, libnccl.so.2): rank 0 calls ptx_comm_unique_id, ships the 128 bytes to
 * the other ranks by any means (MPI, a file, torch.distributed ...), then every rank calls ptx_comm_create -- collectively.
 * A process that also loads another copy of NCCL under the same soname (PyTorch bundles its own libnccl.so.2) must load THAT first
 * (import torch before the first ptx_comm_* call) or point PTX_NCCL_LIB at it: the dynamic loader keeps one libnccl.so.2 per process. */
typedef struct ptx_comm ptx_comm;
#define PTX_COMM_ID_BYTES 128
int ptx_comm_unique_id(uint8_t* id_out, size_t bytes);
int ptx_comm_create(ptx_ctx* ctx, const uint8_t* id, int rank, int world, ptx_comm** out);
void ptx_comm_destroy(ptx_comm* comm);
int ptx_film_combine(ptx_film* film, ptx_comm* comm, int shard_mode);

/* ---- queries used by parity tests (spec: pick_primitive_kernel, ui/callbacks.h:22-48) --------- */
/* pixel-centre primary rays, u=(x+0.5)/w, v=(y+0.5)/h; prim_id in the reference's numbering, -1 = miss */
int ptx_trace_primary_ids(ptx_ctx* ctx, const ptx_scene* scene, const ptx_camera* cam, int width, int height,
                          int32_t* prim_id_host, float* t_host);
/* closest hit for n host rays (6 floats: o, d; d normalised like the Ray ctor, core/ray.h:9-12) */
int ptx_trace_rays(ptx_ctx* ctx, const ptx_scene* scene, const float* rays_host, size_t n, float t_min,
                   int32_t* prim_id_host, float* t_host);
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and results as
 * ptx_trace_rays, but the call allocates nothing: the device staging for rays / ids / distances is a file-scope device array
 * of the library (2 MB, held by every GPU that loads the library until exit), so n is limited to PTX_NOALLOC_MAX_RAYS per call
 * (more is rejected with PTX_ERR_INVALID).  All calls share that array: it keeps the last call's contents (never read beyond
 * the current n), and calls that overlap on the GPU -- two contexts on one device, two host threads -- race on it. */
#define PTX_NOALLOC_MAX_RAYS 65536
int ptx_trace_rays_noalloc(ptx_ctx* ctx, const ptx_scene* scene, const float* rays_host, size_t n, float t_min,
                           int32_t* prim_id_host, float* t_host);
/* any-hit shadow query: Scene::occluded(from, to) (scene.h:132-190) for n segments of 6 floats (from.xyz, to.xyz);
 * occluded_host[i] = 1 if something lies strictly between the two points (1e-4 off both ends, as the reference) */
int ptx_trace_occluded(ptx_ctx* ctx, const ptx_scene* scene, const float* segs_host, size_t n, uint8_t* occluded_host);
/* primary-visibility view: the render_radiosity kernel (integrator.h:460-504).  spp jittered primary rays per pixel from
 * the film's RNG streams (advanced like the reference does); colour = Le + radiosity_host[3 * primitive] (3 floats per
 * reference primitive, NULL = 0: the radiosity solver that fills them is outside this library), sqrt, clamp, RGB8
 * [h][w][3] with row 0 at the bottom */
int ptx_render_visibility(ptx_ctx* ctx, const ptx_scene* scene, const ptx_camera* cam, ptx_film* film, int spp,
                          const float* radiosity_host, uint8_t* rgb8_host);
/* XORWOW state after render_init for pixels [first, first+n): 6 words each (v[0..4], d) */
int ptx_film_rng_state(ptx_film* film, int64_t first, int64_t n, uint32_t* out_host);

#ifdef __cplusplus
}
#endif
#endif /* PTX_H */
