// Is the packed fp32 FMA (fma.rn.f32x2 -> FFMA2) twice as fast per element as scalar FFMA on sm_100a?
// Times the degree-17 normal-CDF polynomial of common.cuh both ways.  nvcc -arch=sm_100a -O3 ffma2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void up(u64 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__constant__ float kC[9] = {1.464686745e+00f, -7.833959951e+00f, 1.873559503e+01f, -2.683502153e+01f, 2.613513962e+01f,
                            -1.863370157e+01f, 1.012229878e+01f, -4.250754193e+00f, 1.595699084e+00f};
__device__ __forceinline__ float cdf1(float x) {
    x = fminf(fmaxf(x, -4.f), 4.f);
    const float t = x * 0.25f, t2 = t * t;
    float q = 1.464686745e+00f;
    q = fmaf(q, t2, -7.833959951e+00f); q = fmaf(q, t2, 1.873559503e+01f); q = fmaf(q, t2, -2.683502153e+01f);
    q = fmaf(q, t2, 2.613513962e+01f); q = fmaf(q, t2, -1.863370157e+01f); q = fmaf(q, t2, 1.012229878e+01f);
    q = fmaf(q, t2, -4.250754193e+00f); q = fmaf(q, t2, 1.595699084e+00f);
    return fmaf(q, t, 0.5f);
}
__device__ __forceinline__ void cdf2(float x0, float x1, float& y0, float& y1) {
    x0 = fminf(fmaxf(x0, -4.f), 4.f); x1 = fminf(fmaxf(x1, -4.f), 4.f);
    const u64 t = mul2(pk(x0, x1), pk(0.25f, 0.25f)), t2 = mul2(t, t);
    u64 q = pk(1.464686745e+00f, 1.464686745e+00f);
    q = fma2(q, t2, pk(-7.833959951e+00f, -7.833959951e+00f)); q = fma2(q, t2, pk(1.873559503e+01f, 1.873559503e+01f));
    q = fma2(q, t2, pk(-2.683502153e+01f, -2.683502153e+01f)); q = fma2(q, t2, pk(2.613513962e+01f, 2.613513962e+01f));
    q = fma2(q, t2, pk(-1.863370157e+01f, -1.863370157e+01f)); q = fma2(q, t2, pk(1.012229878e+01f, 1.012229878e+01f));
    q = fma2(q, t2, pk(-4.250754193e+00f, -4.250754193e+00f)); q = fma2(q, t2, pk(1.595699084e+00f, 1.595699084e+00f));
    up(fma2(q, t, pk(0.5f, 0.5f)), y0, y1);
}
template <int MODE>
__global__ void k(float* out, int iters) {
    float x[8];
    for (int j = 0; j < 8; ++j) x[j] = (threadIdx.x * 8 + j) * 1e-3f - 1.f;
    float acc = 0.f;
    for (int i = 0; i < iters; ++i) {
        if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float y = cdf1(x[j]); acc += y; x[j] = x[j] * 0.999f + y * 1e-3f; }
        } else {
#pragma unroll
            for (int j = 0; j < 8; j += 2) { float y0, y1; cdf2(x[j], x[j + 1], y0, y1); acc += y0 + y1; x[j] = x[j] * 0.999f + y0 * 1e-3f; x[j + 1] = x[j + 1] * 0.999f + y1 * 1e-3f; }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int mode = 0; mode < 2; ++mode) for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (mode == 0) k<0><<<148 * 8, 256>>>(d, 4096); else k<1><<<148 * 8, 256>>>(d, 4096);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        const double evals = 148.0 * 8 * 256 * 8 * 4096;
        printf("mode %d (%s): %.3f ms, %.1f G evals/s\n", mode, mode ? "FFMA2" : "FFMA", ms, evals / ms * 1e-6);
    }
    float h[4]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost); printf("%g %s\n", h[1], cudaGetErrorString(cudaGetLastError()));
    return 0;
}
