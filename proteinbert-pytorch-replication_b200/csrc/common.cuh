// Shared device/host helpers for the ProteinBERT B200 kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#define PB_OK 0
#define PB_ERR_SHAPE 1
#define PB_ERR_CUDA 2
#define PB_ERR_UNSUPPORTED 3
#define PB_ERR_WORKSPACE 4

#define PB_F32 0
#define PB_BF16 1

#define PB_ACT_NONE 0
#define PB_ACT_GELU 1
#define PB_ACT_SIGMOID 2

typedef __nv_bfloat16 bf16;

// ---- error plumbing ---------------------------------------------------------
void pb_set_error(const char* fmt, ...);
int pb_check_launch(const char* what);
void pb_count_launch(int n);

#define PB_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            pb_set_error(__VA_ARGS__);         \
            return (code);                     \
        }                                      \
    } while (0)

#define PB_CUDA(call)                                                            \
    do {                                                                         \
        cudaError_t e__ = (call);                                                \
        if (e__ != cudaSuccess) {                                                \
            pb_set_error("%s failed: %s", #call, cudaGetErrorString(e__));       \
            return PB_ERR_CUDA;                                                  \
        }                                                                        \
    } while (0)

// ---- scalar helpers ---------------------------------------------------------
__device__ __forceinline__ float ldf(const float* p, size_t i) { return p[i]; }
__device__ __forceinline__ float ldf(const bf16* p, size_t i) { return __bfloat162float(p[i]); }
__device__ __forceinline__ void stf(float* p, size_t i, float v) { p[i] = v; }
__device__ __forceinline__ void stf(bf16* p, size_t i, float v) { p[i] = __float2bfloat16_rn(v); }
// value after a round trip through the storage type (stats are taken of STORED values)
__device__ __forceinline__ float rnd(const float*, float v) { return v; }
__device__ __forceinline__ float rnd(const bf16*, float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// exact-erf GELU, as nn.GELU() (approximate='none')
__device__ __forceinline__ float gelu_f(float t) { return 0.5f * t * (1.0f + erff(t * 0.70710678118654752440f)); }
// d gelu / dt = Phi(t) + t*phi(t)
__device__ __forceinline__ float gelu_grad_f(float t) {
    const float cdf = 0.5f * (1.0f + erff(t * 0.70710678118654752440f));
    const float pdf = 0.39894228040143267794f * __expf(-0.5f * t * t);
    return cdf + t * pdf;
}
// --- fast variants for the bf16 kernels (results are stored as bf16, eps = 3.9e-3) -------------------------
// Normal CDF by an odd degree-17 polynomial in x on |x| <= 4 (Chebyshev fit in x/4, coefficients rescaled to x;
// |abs error| <= 1.6e-5 in fp32 Horner form): FMA pipe only, no MUFU, 13 instructions.  Above 4 the argument is
// clamped (the value there is 1 - 3e-5), below -4 the result is exactly 0 so that sums of many strongly negative
// pre-activations (the attention pooling) collect no bias.  max |gelu error| 6.3e-5 on [-4, 4].
__device__ __forceinline__ float normal_cdf_fast(float x) {
    const float xc = fminf(x, 4.f), x2 = xc * xc;
    float q = 8.525598940e-11f;
    q = fmaf(q, x2, -7.295943751e-09f);
    q = fmaf(q, x2, 2.791821216e-07f);
    q = fmaf(q, x2, -6.397967703e-06f);
    q = fmaf(q, x2, 9.969764565e-05f);
    q = fmaf(q, x2, -1.137310887e-03f);
    q = fmaf(q, x2, 9.885057402e-03f);
    q = fmaf(q, x2, -6.641803427e-02f);
    q = fmaf(q, x2, 3.989247710e-01f);
    const float c = fmaf(q, xc, 0.5f);
    return x <= -4.f ? 0.f : c;
}
__device__ __forceinline__ float gelu_fast(float t) { return t * normal_cdf_fast(t); }
// gelu'(x) = Phi(x) + x phi(x) directly: gelu'(x) - 1/2 is odd, fitted by a degree-19 odd polynomial on |x| <= 4 (weighted
// Chebyshev fit, |abs error| <= 4.4e-5 in fp32 Horner form), argument clamped outside (there gelu' is within 5e-4 of its
// value at +-4: 1.0005 / -0.0005).  13 FP32 instructions, no MUFU; the previous form (Phi polynomial + ex2) took 19 and
// the kernels that call this are bound by the FP32 pipe.
__device__ __forceinline__ float gelu_grad_fast(float x) {
    const float xc = fminf(fmaxf(x, -4.f), 4.f), x2 = xc * xc;
    float q = -5.066656453e-11f;
    q = fmaf(q, x2, 4.789404823e-09f);
    q = fmaf(q, x2, -2.026334901e-07f);
    q = fmaf(q, x2, 5.107091856e-06f);
    q = fmaf(q, x2, -8.623214197e-05f);
    q = fmaf(q, x2, 1.038562907e-03f);
    q = fmaf(q, x2, -9.190203826e-03f);
    q = fmaf(q, x2, 5.937874753e-02f);
    q = fmaf(q, x2, -2.656380801e-01f);
    q = fmaf(q, x2, 7.978171689e-01f);
    return fmaf(q, xc, 0.5f);
}

// --- packed pairs: fma.rn.f32x2 / mul.rn.f32x2 (SASS FFMA2 / FMUL2) evaluate two independent fp32 operations per issued
// instruction, each rounded exactly like the scalar one.  The FP32 pipe does the same work (a microbenchmark that saturates the
// pipe gains nothing: profiles/r01_ffma2_probe.md), but the GELU epilogues of this library are bound by instruction ISSUE
// (ncu: issue slots 62-68 % busy, FMA pipe ~40 %), and the polynomial is 10 of their ~15-25 instructions per element.
typedef unsigned long long pb_f32x2;
__device__ __forceinline__ pb_f32x2 pk2(float a, float b) { pb_f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void up2(pb_f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ pb_f32x2 fma2(pb_f32x2 a, pb_f32x2 b, pb_f32x2 c) {
    pb_f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;
}
__device__ __forceinline__ pb_f32x2 mul2(pb_f32x2 a, pb_f32x2 b) { pb_f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
#define PB_BOTH(c) pk2(c, c)
// Phi(x) for two arguments, argument clamped to [-4, 4] on both sides.  The constant term is 0.5 - Phi_poly(-4) = 0.4999826f,
// which makes the value at and below -4 EXACTLY 0 in fp32 (checked by evaluating the Horner form with single-rounding FMAs),
// as normal_cdf_fast's select does, without the compare + select; everywhere else Phi moves by -1.74e-5 (the polynomial's own
// error is 1.6e-5).  max |gelu error| 2.1e-4 (at x = 6; 1.3e-4 for gelu_fast), far inside bf16 rounding.
__device__ __forceinline__ pb_f32x2 normal_cdf_fast2(float x0, float x1, pb_f32x2& xc) {
    xc = pk2(fminf(fmaxf(x0, -4.f), 4.f), fminf(fmaxf(x1, -4.f), 4.f));
    const pb_f32x2 x2 = mul2(xc, xc);
    pb_f32x2 q = fma2(PB_BOTH(8.525598940e-11f), x2, PB_BOTH(-7.295943751e-09f));
    q = fma2(q, x2, PB_BOTH(2.791821216e-07f));
    q = fma2(q, x2, PB_BOTH(-6.397967703e-06f));
    q = fma2(q, x2, PB_BOTH(9.969764565e-05f));
    q = fma2(q, x2, PB_BOTH(-1.137310887e-03f));
    q = fma2(q, x2, PB_BOTH(9.885057402e-03f));
    q = fma2(q, x2, PB_BOTH(-6.641803427e-02f));
    q = fma2(q, x2, PB_BOTH(3.989247710e-01f));
    return fma2(q, xc, PB_BOTH(0.4999826f));
}
// y = gelu(t) for two values
__device__ __forceinline__ void gelu_fast2(float t0, float t1, float& y0, float& y1) {
    pb_f32x2 xc;
    const pb_f32x2 c = normal_cdf_fast2(t0, t1, xc);
    up2(mul2(pk2(t0, t1), c), y0, y1);
}
// y = gelu'(t) for two values (the polynomial of gelu_grad_fast)
__device__ __forceinline__ void gelu_grad_fast2(float t0, float t1, float& y0, float& y1) {
    const pb_f32x2 xc = pk2(fminf(fmaxf(t0, -4.f), 4.f), fminf(fmaxf(t1, -4.f), 4.f));
    const pb_f32x2 x2 = mul2(xc, xc);
    pb_f32x2 q = fma2(PB_BOTH(-5.066656453e-11f), x2, PB_BOTH(4.789404823e-09f));
    q = fma2(q, x2, PB_BOTH(-2.026334901e-07f));
    q = fma2(q, x2, PB_BOTH(5.107091856e-06f));
    q = fma2(q, x2, PB_BOTH(-8.623214197e-05f));
    q = fma2(q, x2, PB_BOTH(1.038562907e-03f));
    q = fma2(q, x2, PB_BOTH(-9.190203826e-03f));
    q = fma2(q, x2, PB_BOTH(5.937874753e-02f));
    q = fma2(q, x2, PB_BOTH(-2.656380801e-01f));
    q = fma2(q, x2, PB_BOTH(7.978171689e-01f));
    up2(fma2(q, xc, PB_BOTH(0.5f)), y0, y1);
}
// in place over an array of even length
template <int N>
__device__ __forceinline__ void gelu_fast_n(float (&v)[N]) {
#pragma unroll
    for (int j = 0; j < N; j += 2) gelu_fast2(v[j], v[j + 1], v[j], v[j + 1]);
}
template <int N>
__device__ __forceinline__ void gelu_grad_fast_n(float (&v)[N]) {
#pragma unroll
    for (int j = 0; j < N; j += 2) gelu_grad_fast2(v[j], v[j + 1], v[j], v[j + 1]);
}

__device__ __forceinline__ float sigmoid_f(float t) { return 1.0f / (1.0f + expf(-t)); }

__device__ __forceinline__ float act_fwd(int act, float z) {
    return act == PB_ACT_GELU ? gelu_f(z) : (act == PB_ACT_SIGMOID ? sigmoid_f(z) : z);
}
// derivative given pre-activation z (gelu) or OUTPUT y (sigmoid): the caller passes the right one
__device__ __forceinline__ float act_bwd(int act, float zy) {
    return act == PB_ACT_GELU ? gelu_grad_f(zy) : (act == PB_ACT_SIGMOID ? zy * (1.0f - zy) : 1.0f);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// block-wide sum for blockDim.x <= 1024; result valid in every thread
template <typename V>
__device__ __forceinline__ V block_sum(V v, V* smem /* >= 32 */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) smem[w] = v;
    __syncthreads();
    V r = (threadIdx.x < nw) ? smem[threadIdx.x] : V(0);
    if (w == 0) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
        if (lane == 0) smem[0] = r;
    }
    __syncthreads();
    return smem[0];
}

// per-sample LayerNorm statistics from the (sum, sumsq) doubles
struct LnStat { float mean, rstd; };
__device__ __forceinline__ LnStat ln_stat(const double* stats, int b, double n, float eps) {
    const double s = stats[2 * b], q = stats[2 * b + 1];
    const double mean = s / n;
    double var = q / n - mean * mean;
    if (var < 0.0) var = 0.0;
    LnStat r;
    r.mean = (float)mean;
    r.rstd = (float)(1.0 / sqrt(var + (double)eps));
    return r;
}

static inline int ceil_div(long a, long b) { return (int)((a + b - 1) / b); }

// ---- programmatic dependent launch ----------------------------------------------------------------------
// Kernels launched through pb_launch_pdl may be scheduled while the previous kernel in the stream drains;
// they call pdl_wait() before touching global memory, which blocks until that kernel has completed and its
// writes are visible.  This hides the launch latency between the ~200 short kernels of one step.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
bool pb_pdl_enabled();

template <typename... KArgs, typename... Args>
static inline cudaError_t pb_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pb_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
