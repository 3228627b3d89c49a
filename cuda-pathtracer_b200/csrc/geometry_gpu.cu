// geometry_gpu.cu -- see geometry_gpu.h
#include "geometry_gpu.h"

#include <cub/cub.cuh>

namespace ptx {

namespace {

constexpr int kT = 256;
inline unsigned grid_for(size_t n) { return (unsigned)((n + kT - 1) / kT); }

#define GG_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { err = std::string(#call) + ": " + cudaGetErrorString(e_); return e_; } } while (0)

__device__ __forceinline__ void set3(float* d, vec3 a) { d[0] = a.x; d[1] = a.y; d[2] = a.z; }
// Geometric normal of the Triangle / Quad ctor (triangle.h:21-29, quad.h:23-31).  The reference runs this on the HOST,
// where the compiler does not contract a*b - c*d into an FMA; the device code must not either, or the normals (and with
// them every shaded sample) would differ in the last bit from a host-preprocessed scene: explicit _rn intrinsics.
__device__ __forceinline__ void geo_normal(const float* a, const float* b, const float* c, float* n) {
    vec3 p = ld3(b) - ld3(a), q = ld3(c) - ld3(a);
    float cx = __fsub_rn(__fmul_rn(p.y, q.z), __fmul_rn(p.z, q.y));
    float cy = -__fsub_rn(__fmul_rn(p.x, q.z), __fmul_rn(p.z, q.x));
    float cz = __fsub_rn(__fmul_rn(p.x, q.y), __fmul_rn(p.y, q.x));
    float l2 = __fadd_rn(__fadd_rn(__fmul_rn(cx, cx), __fmul_rn(cy, cy)), __fmul_rn(cz, cz));
    float k = __frcp_rn(__fsqrt_rn(l2));
    n[0] = __fmul_rn(cx, k); n[1] = __fmul_rn(cy, k); n[2] = __fmul_rn(cz, k);
}
__device__ __forceinline__ vec3 mid(const float* a, const float* b) { return (ld3(a) + ld3(b)) * 0.5f; }

// 1 output per triangle, 2 per quad (when splitting) -- also the triangle count per primitive for the packer
__global__ void __launch_bounds__(kT) k_count(const ptx_primitive* __restrict__ in, size_t n, uint32_t* cnt) {
    size_t i = (size_t)blockIdx.x * kT + threadIdx.x;
    if (i < n) cnt[i] = in[i].type == 1 ? 2u : 1u;
}

// convertQuadsToTriangles: (v00,v10,v11), (v00,v11,v01), geometric normals, material carried over
__global__ void __launch_bounds__(kT) k_split_quads(const ptx_primitive* __restrict__ in, const uint32_t* __restrict__ in_mat, size_t n,
                                                    const uint32_t* __restrict__ offs, ptx_primitive* out, uint32_t* out_mat) {
    size_t i = (size_t)blockIdx.x * kT + threadIdx.x;
    if (i >= n) return;
    ptx_primitive p = in[i];
    uint32_t o = offs[i], m = in_mat[i];
    if (p.type == 0) { out[o] = p; out_mat[o] = m; return; }
    for (int h = 0; h < 2; h++) {
        ptx_primitive t = p;
        t.type = 0;
        for (int k = 0; k < 3; k++) { t.v[1][k] = p.v[1 + h][k]; t.v[2][k] = p.v[2 + h][k]; t.v[3][k] = 0.f; }
        geo_normal(t.v[0], t.v[1], t.v[2], t.normal);
        out[o + h] = t; out_mat[o + h] = m;
    }
}

// one subdivision level: primitive i -> outputs 4i .. 4i+3 (corner, corner, corner, centre / four quads round the centre)
__global__ void __launch_bounds__(kT) k_subdivide(const ptx_primitive* __restrict__ in, const uint32_t* __restrict__ in_mat, size_t n,
                                                  ptx_primitive* out, uint32_t* out_mat) {
    size_t i = (size_t)blockIdx.x * kT + threadIdx.x;
    if (i >= n) return;
    ptx_primitive p = in[i];
    uint32_t m = in_mat[i];
    if (p.type == 0) {
        float c[6][3];                                         // v0 v1 v2 m0 m1 m2
        for (int k = 0; k < 3; k++) { c[0][k] = p.v[0][k]; c[1][k] = p.v[1][k]; c[2][k] = p.v[2][k]; }
        set3(c[3], mid(p.v[0], p.v[1])); set3(c[4], mid(p.v[1], p.v[2])); set3(c[5], mid(p.v[2], p.v[0]));
        const int idx[4][3] = {{0, 3, 5}, {3, 1, 4}, {4, 2, 5}, {3, 4, 5}};
        for (int s = 0; s < 4; s++) {
            ptx_primitive t = p;
            for (int k = 0; k < 3; k++) for (int a = 0; a < 3; a++) t.v[k][a] = c[idx[s][k]][a];
            geo_normal(t.v[0], t.v[1], t.v[2], t.normal);
            out[4 * i + s] = t; out_mat[4 * i + s] = m;
        }
    } else {
        float c[9][3];                                         // v0..v3, m01 m12 m23 m30, centre
        for (int q = 0; q < 4; q++) for (int k = 0; k < 3; k++) c[q][k] = p.v[q][k];
        set3(c[4], mid(p.v[0], p.v[1])); set3(c[5], mid(p.v[1], p.v[2])); set3(c[6], mid(p.v[2], p.v[3])); set3(c[7], mid(p.v[3], p.v[0]));
        set3(c[8], (((ld3(p.v[0]) + ld3(p.v[1])) + ld3(p.v[2])) + ld3(p.v[3])) * 0.25f);
        const int idx[4][4] = {{0, 4, 8, 7}, {4, 1, 5, 8}, {8, 5, 2, 6}, {7, 8, 6, 3}};
        for (int s = 0; s < 4; s++) {
            ptx_primitive t = p;
            for (int k = 0; k < 4; k++) for (int a = 0; a < 3; a++) t.v[k][a] = c[idx[s][k]][a];
            geo_normal(t.v[0], t.v[1], t.v[3], t.normal);
            out[4 * i + s] = t; out_mat[4 * i + s] = m;
        }
    }
}

// builder input + shading records (scene_pack.cpp: pack_primitives)
__global__ void __launch_bounds__(kT) k_pack(const ptx_primitive* __restrict__ in, const uint32_t* __restrict__ in_mat, size_t n,
                                             const uint32_t* __restrict__ offs, TriIn* tris, f32x4* prim_shade) {
    size_t i = (size_t)blockIdx.x * kT + threadIdx.x;
    if (i >= n) return;
    ptx_primitive p = in[i];
    prim_shade[i] = f32x4{p.normal[0], p.normal[1], p.normal[2], u2f(in_mat[i])};
    TriIn t;
    for (int k = 0; k < 3; k++) { t.v0[k] = p.v[0][k]; t.v1[k] = p.v[1][k]; t.v2[k] = p.v[2][k]; }
    t.prim_id = (uint32_t)i;
    uint32_t o = offs[i];
    tris[o] = t;
    if (p.type == 1) {
        for (int k = 0; k < 3; k++) { t.v1[k] = p.v[2][k]; t.v2[k] = p.v[3][k]; }
        tris[o + 1] = t;
    }
}

// counts -> exclusive offsets (in place into `offs`), total returned on the host
cudaError_t scan_counts(cudaStream_t st, const ptx_primitive* d_prims, size_t n, uint32_t* d_cnt, uint32_t* d_offs, size_t& total, std::string& err) {
    k_count<<<grid_for(n), kT, 0, st>>>(d_prims, n, d_cnt);
    size_t tmp_bytes = 0;
    GG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_cnt, d_offs, (int)n, st));
    void* d_tmp = nullptr;
    GG_CUDA(cudaMalloc(&d_tmp, tmp_bytes ? tmp_bytes : 16));
    cudaError_t e = cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_cnt, d_offs, (int)n, st);
    uint32_t last_off = 0, last_cnt = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&last_off, d_offs + n - 1, 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&last_cnt, d_cnt + n - 1, 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d_tmp);
    if (e != cudaSuccess) { err = std::string("scan_counts: ") + cudaGetErrorString(e); return e; }
    total = (size_t)last_off + last_cnt;
    return cudaSuccess;
}

}  // namespace

void DevGeometry::release() {
    cudaFree(d_prims); cudaFree(d_mat); cudaFree(d_tris); cudaFree(d_prim_shade);
    d_prims = nullptr; d_mat = nullptr; d_tris = nullptr; d_prim_shade = nullptr; n_prims = n_tris = 0;
}

cudaError_t preprocess_geometry_gpu(cudaStream_t st, const ptx_primitive* h_base, const uint32_t* h_base_mat, size_t n0,
                                    bool convert_quads, int levels, DevGeometry& out, std::string& err) {
    out = DevGeometry();
    if (n0 == 0) return cudaSuccess;
    if (levels < 0) levels = 0;
    size_t n_max = n0 * (convert_quads ? 2 : 1);
    for (int l = 0; l < levels; l++) {
        if (n_max > (size_t)0x7FFFFFF0u / 8) { err = "subdivision would exceed 2^31 triangles"; return cudaErrorInvalidValue; }
        n_max *= 4;
    }
    cudaEvent_t e0, e1;
    GG_CUDA(cudaEventCreate(&e0)); GG_CUDA(cudaEventCreate(&e1));
    ptx_primitive *cur = nullptr, *nxt = nullptr; uint32_t *cur_m = nullptr, *nxt_m = nullptr, *d_cnt = nullptr, *d_offs = nullptr;
    auto cleanup = [&]() { cudaFree(nxt); cudaFree(nxt_m); cudaFree(d_cnt); cudaFree(d_offs); cudaEventDestroy(e0); cudaEventDestroy(e1); };
    cudaError_t e = cudaMalloc(&cur, sizeof(ptx_primitive) * n0);
    if (e == cudaSuccess) e = cudaMalloc(&cur_m, 4 * n0);
    if (e == cudaSuccess) e = cudaMemcpyAsync(cur, h_base, sizeof(ptx_primitive) * n0, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(cur_m, h_base_mat, 4 * n0, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaEventRecord(e0, st);
    size_t n = n0;
    if (e == cudaSuccess && convert_quads) {
        e = cudaMalloc(&d_cnt, 4 * n); if (e == cudaSuccess) e = cudaMalloc(&d_offs, 4 * n);
        size_t total = 0;
        if (e == cudaSuccess) e = scan_counts(st, cur, n, d_cnt, d_offs, total, err);
        if (e == cudaSuccess) e = cudaMalloc(&nxt, sizeof(ptx_primitive) * total);
        if (e == cudaSuccess) e = cudaMalloc(&nxt_m, 4 * total);
        if (e == cudaSuccess) {
            k_split_quads<<<grid_for(n), kT, 0, st>>>(cur, cur_m, n, d_offs, nxt, nxt_m);
            e = cudaStreamSynchronize(st);                     // the inputs are freed next
        }
        if (e == cudaSuccess) { cudaFree(cur); cudaFree(cur_m); cur = nxt; cur_m = nxt_m; nxt = nullptr; nxt_m = nullptr; n = total; }
        cudaFree(d_cnt); cudaFree(d_offs); d_cnt = d_offs = nullptr;
    }
    for (int l = 0; l < levels && e == cudaSuccess; l++) {
        e = cudaMalloc(&nxt, sizeof(ptx_primitive) * n * 4);
        if (e == cudaSuccess) e = cudaMalloc(&nxt_m, 4 * n * 4);
        if (e == cudaSuccess) {
            k_subdivide<<<grid_for(n), kT, 0, st>>>(cur, cur_m, n, nxt, nxt_m);
            e = cudaStreamSynchronize(st);
        }
        if (e == cudaSuccess) { cudaFree(cur); cudaFree(cur_m); cur = nxt; cur_m = nxt_m; nxt = nullptr; nxt_m = nullptr; n *= 4; }
    }
    TriIn* d_tris = nullptr; f32x4* d_ps = nullptr; size_t n_tris = 0;
    if (e == cudaSuccess) {
        e = cudaMalloc(&d_cnt, 4 * n); if (e == cudaSuccess) e = cudaMalloc(&d_offs, 4 * n);
        if (e == cudaSuccess) e = scan_counts(st, cur, n, d_cnt, d_offs, n_tris, err);
        if (e == cudaSuccess) e = cudaMalloc(&d_tris, sizeof(TriIn) * n_tris);
        if (e == cudaSuccess) e = cudaMalloc(&d_ps, sizeof(f32x4) * n);
        if (e == cudaSuccess) {
            k_pack<<<grid_for(n), kT, 0, st>>>(cur, cur_m, n, d_offs, d_tris, d_ps);
            e = cudaEventRecord(e1, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
    }
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
        if (err.empty()) err = std::string("preprocess_geometry_gpu: ") + cudaGetErrorString(e);
        cudaFree(cur); cudaFree(cur_m); cudaFree(d_tris); cudaFree(d_ps); cleanup();
        return e;
    }
    cudaEventElapsedTime(&out.ms, e0, e1);
    out.d_prims = cur; out.d_mat = cur_m; out.n_prims = n; out.d_tris = d_tris; out.n_tris = n_tris; out.d_prim_shade = d_ps;
    cleanup();
    return cudaSuccess;
}

}  // namespace ptx
