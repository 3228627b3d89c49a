// tools/ubench/exact_math.cu -- exhaustive check of the branch-free "exact" reciprocal / square root / division used by the
// kernels (csrc/common.h) against CUDA's correctly rounded intrinsics, over EVERY float32 bit pattern.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I cuda-pathtracer_b200/csrc -o /tmp/exact_math tools/ubench/exact_math.cu && /tmp/exact_math
// Prints, per function, the number of inputs inside the documented domain and how many of them differ (must be 0), and
// what happens outside the domain.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "common.h"

using namespace ptx;

__device__ bool same(float a, float b) { return __float_as_uint(a) == __float_as_uint(b) || (a != a && b != b); }

// counters: [0] in-domain inputs, [1] in-domain mismatches, [2] out-of-domain inputs, [3] out-of-domain mismatches
__global__ void k_rcp(unsigned long long* c) {
    unsigned long long in = 0, bad = 0, out = 0, obad = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < (1ull << 32); i += (unsigned long long)gridDim.x * blockDim.x) {
        float x = __uint_as_float((uint32_t)i);
        uint32_t e = ((uint32_t)i >> 23) & 0xFFu;
        bool dom = e >= 1u && e <= 252u;                        // 2^-126 <= |x| < 2^126
        bool ok = same(rcp_exact(x), __frcp_rn(x));
        if (dom) { in++; bad += !ok; } else { out++; obad += !ok; }
    }
    atomicAdd(c + 0, in); atomicAdd(c + 1, bad); atomicAdd(c + 2, out); atomicAdd(c + 3, obad);
}
__global__ void k_sqrt(unsigned long long* c) {
    unsigned long long in = 0, bad = 0, out = 0, obad = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < (1ull << 32); i += (unsigned long long)gridDim.x * blockDim.x) {
        float x = __uint_as_float((uint32_t)i);
        bool dom = (uint32_t)i == 0u || ((uint32_t)i >= 0x0D000000u && (uint32_t)i <= 0x7F7FFFFFu);   // +0 and [2^-101, FLT_MAX]
        bool ok = same(sqrt_exact(x), __fsqrt_rn(x));
        if (dom) { in++; bad += !ok; } else { out++; obad += !ok; }
    }
    atomicAdd(c + 0, in); atomicAdd(c + 1, bad); atomicAdd(c + 2, out); atomicAdd(c + 3, obad);
}
// division a / b for the camera jitter (start_path_xy): b = an image dimension, a = pixel coordinate + jitter in [0, 65536]
__global__ void k_div(unsigned long long* c, float b) {
    unsigned long long in = 0, bad = 0, out = 0, obad = 0;
    const uint32_t lo = __float_as_uint(9.094947017729282e-13f), hi = __float_as_uint(65536.0f);      // 2^-40 .. 2^16
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i <= hi; i += (unsigned long long)gridDim.x * blockDim.x) {
        float a = __uint_as_float((uint32_t)i);                  // every float in [0, 65536]
        bool ok = same(div_exact_pos(a, b), __fdiv_rn(a, b));
        if ((uint32_t)i >= lo) { in++; bad += !ok; } else { out++; obad += !ok; }
    }
    atomicAdd(c + 0, in); atomicAdd(c + 1, bad); atomicAdd(c + 2, out); atomicAdd(c + 3, obad);
}

// sincos_small against sincosf for every float32 in [0, 8] (the azimuth 2 pi v lies in (0, 2 pi])
__global__ void k_sincos(unsigned long long* c) {
    unsigned long long in = 0, bad = 0;
    const uint32_t hi = __float_as_uint(8.0f);
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i <= hi; i += (unsigned long long)gridDim.x * blockDim.x) {
        float x = __uint_as_float((uint32_t)i), s0, c0, s1, c1;
        sincosf(x, &s0, &c0);
        sincos_small(x, s1, c1);
        in++; bad += !(same(s0, s1) && same(c0, c1));
    }
    atomicAdd(c + 0, in); atomicAdd(c + 1, bad);
}

// EVERY divisor 1 ... 65535 (ptx_film_create accepts any image dimension up to 65535): all 2^23 mantissas of the numerator at three
// exponents (2^-33: the smallest jitter, 2^0, and 2^15: the largest coordinate range).  The sequence of div_exact_pos is scale-invariant
// in the numerator (r depends on b only; a r, a - b q and q + r e are IEEE operations on scaled operands, nothing under- or overflows
// for 2^-40 <= a <= 65536), so a mantissa that divides exactly at one exponent does at every exponent of the domain; the three
// exponents and the exhaustive all-exponent pass over 35 dimensions below check that claim as well.
__global__ void k_div_all(unsigned long long* c) {
    const float b = (float)(blockIdx.y + 1);
    unsigned long long bad = 0;
    const uint32_t chunk = (1u << 23) / gridDim.x;
    for (uint32_t m = blockIdx.x * chunk + threadIdx.x; m < (blockIdx.x + 1) * chunk; m += blockDim.x) {
        const float a0 = __uint_as_float((94u << 23) | m), a1 = __uint_as_float((127u << 23) | m), a2 = __uint_as_float((142u << 23) | m);
        bad += !same(div_exact_pos(a0, b), __fdiv_rn(a0, b));
        bad += !same(div_exact_pos(a1, b), __fdiv_rn(a1, b));
        bad += !same(div_exact_pos(a2, b), __fdiv_rn(a2, b));
    }
    bad = __reduce_add_sync(0xFFFFFFFFu, (unsigned)bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(c + 1, bad);
    if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(c + 0, 3ull << 23);
}

int main() {
    unsigned long long* d; cudaMalloc(&d, 64);
    unsigned long long h[4];
    auto run = [&](const char* name, auto launch) {
        cudaMemset(d, 0, 64); launch(); cudaDeviceSynchronize();
        cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
        printf("%-28s in-domain %llu mismatches %llu | out-of-domain %llu mismatches %llu\n", name, h[0], h[1], h[2], h[3]);
        return h[1];
    };
    unsigned long long fails = 0;
    fails += run("rcp_exact vs __frcp_rn", [&] { k_rcp<<<148 * 8, 256>>>(d); });
    fails += run("sqrt_exact vs __fsqrt_rn", [&] { k_sqrt<<<148 * 8, 256>>>(d); });
    const int dims[] = {1, 2, 3, 7, 8, 36, 37, 40, 48, 54, 64, 91, 96, 100, 256, 480, 512, 640, 720, 800, 1000, 1024, 1080, 1085, 1200, 1920, 2048, 2160, 3840, 4096, 7680, 8192, 16384, 32768, 65535};
    unsigned long long tot = 0, bad = 0, otot = 0, obad = 0;
    for (int w : dims) {
        cudaMemset(d, 0, 64); k_div<<<148 * 8, 256>>>(d, (float)w); cudaDeviceSynchronize();
        cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost); tot += h[0]; bad += h[1]; otot += h[2]; obad += h[3];
    }
    printf("%-28s in-domain %llu mismatches %llu | out-of-domain %llu mismatches %llu (a: every float in [2^-40, 65536] / below; b: %zu image dimensions)\n",
           "div_exact_pos vs __fdiv_rn", tot, bad, otot, obad, sizeof dims / sizeof dims[0]);
    fails += bad;
    cudaMemset(d, 0, 64); k_div_all<<<dim3(32, 65535), 256>>>(d); cudaDeviceSynchronize();
    cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
    printf("%-28s checks %llu mismatches %llu (EVERY divisor 1 ... 65535 x every numerator mantissa x exponents 2^-33, 2^0, 2^15)\n", "div_exact_pos, all divisors", h[0], h[1]);
    fails += h[1];
    cudaMemset(d, 0, 64); k_sincos<<<148 * 8, 256>>>(d); cudaDeviceSynchronize();
    cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
    printf("%-28s in-domain %llu mismatches %llu | (every float in [0, 8], both outputs)\n", "sincos_small vs sincosf", h[0], h[1]);
    fails += h[1];
    printf(fails ? "FAILED\n" : "ALL EXACT\n");
    return fails ? 1 : 0;
}
