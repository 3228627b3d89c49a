// Issue-slot cost of packed FP32 (FFMA2) against scalar FFMA on sm_100a, alone and mixed with ALU-pipe work.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu && ./ffma2
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters, float s) {
    float a[8]; float2 p[8]; unsigned u[8];
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 0.001f + i; p[i] = make_float2(a[i], a[i] + 0.5f); u[i] = threadIdx.x + i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0 || MODE == 2) { a[i] = fmaf(a[i], s, 0.25f); }
            if (MODE == 4) { a[i] = fmaf(a[i], s, 0.25f); p[i].x = fmaf(p[i].x, s, 0.5f); }   // 16 scalar FMAs = the flops of mode 1
            if (MODE == 1 || MODE == 3) { p[i] = __ffma2_rn(p[i], make_float2(s, s), make_float2(0.25f, 0.5f)); }
            if (MODE == 2 || MODE == 3) { u[i] = (u[i] ^ (u[i] >> 3)) + 0x9E3779B9u; }       // LOP3/SHF + IADD on the ALU pipe
        }
    }
    float r = 0; for (int i = 0; i < 8; i++) r += a[i] + p[i].x + p[i].y + (float)u[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> float run(float* d, int iters) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 8, 256>>>(d, 16, 1.0001f); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<MODE><<<148 * 8, 256>>>(d, iters, 1.0001f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    int iters = 1 << 16;
    const char* names[5] = {"8 FFMA", "8 FFMA2", "8 FFMA + 8x(xor-shift,add)", "8 FFMA2 + 8x(xor-shift,add)", "16 FFMA"};
    float ms[5] = {run<0>(d, iters), run<1>(d, iters), run<2>(d, iters), run<3>(d, iters), run<4>(d, iters)};
    for (int m = 0; m < 5; m++) printf("%-32s %8.3f ms  (%.2f x of 8 FFMA)\n", names[m], ms[m], ms[m] / ms[0]);
    return 0;
}
