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
// The same polynomial with a symmetric clamp (one instruction less: no select).  Below -4 the CDF stays at its value at -4
// (3.2e-5) instead of 0, i.e. gelu(t) = 3.2e-5 t there (true value ~0): an absolute error of 3e-4 at t = -10.  For epilogues
// whose result is ADDED to O(1) terms and rounded to bf16 (conv pair, local linear); the pooling sums keep gelu_fast.
__device__ __forceinline__ float gelu_fast_sym(float t) {
    const float xc = fminf(fmaxf(t, -4.f), 4.f), x2 = xc * xc;
    float q = 8.525598940e-11f;
    q = fmaf(q, x2, -7.295943751e-09f);
    q = fmaf(q, x2, 2.791821216e-07f);
    q = fmaf(q, x2, -6.397967703e-06f);
    q = fmaf(q, x2, 9.969764565e-05f);
    q = fmaf(q, x2, -1.137310887e-03f);
    q = fmaf(q, x2, 9.885057402e-03f);
    q = fmaf(q, x2, -6.641803427e-02f);
    q = fmaf(q, x2, 3.989247710e-01f);
    return t * fmaf(q, xc, 0.5f);
}
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
