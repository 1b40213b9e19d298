// Local-track operators, exact-arithmetic (fp32 SIMT) engine, templated on the activation storage
// type.  In PB_BF16 mode the large contractions are taken over by the tcgen05 kernels in tc_*.cu when
// the shape qualifies (see pb_tc_*); everything here stays as the parity engine and as the engine for
// shapes the tensor-core tiling does not cover.
// Reference call sites: modules.py:124-164 (layers), :205-217 (block forward), :49-60,:87-92 (attention).
#include <string.h>
#include "gemm_simt.cuh"
#include "../../include/pb_b200.h"
#include "tc_api.cuh"

namespace {

#define DISPATCH_DT(dt, ...)                                                   \
    do {                                                                       \
        if ((dt) == PB_F32) { using T = float; __VA_ARGS__; }                  \
        else if ((dt) == PB_BF16) { using T = bf16; __VA_ARGS__; }             \
        else { pb_set_error("unknown dtype code %d", (dt)); return PB_ERR_UNSUPPORTED; } \
    } while (0)

// ---------------------------------------------------------------------------------------------
// loaders / epilogues
// ---------------------------------------------------------------------------------------------
// im2col view of a channels-last sequence: A(m=(b,l), k=(t,c)) = x[b, l + sign*(t-half)*dil, c] or 0
template <typename T>
struct ConvA {
    static constexpr bool kContig = true;
    const T* x; int L, C, dil, half, sign;
    __device__ float operator()(int m, int k, int) const {
        const int t = k / C, c = k - t * C;
        const int b = m / L, l = m - b * L;
        const int l2 = l + sign * (t - half) * dil;
        return (l2 >= 0 && l2 < L) ? ldf(x, ((size_t)b * L + l2) * C + c) : 0.f;
    }
};
// tap-major weight pack [t][n][kc]: B(k=(t,kc), n)
template <typename T>
struct ConvPackB {
    static constexpr bool kContig = true;
    const T* w; int C;
    __device__ float operator()(int k, int n, int) const {
        const int t = k / C, kc = k - t * C;
        return ldf(w, ((size_t)t * C + n) * C + kc);
    }
};
template <typename T>
struct EpBiasStore {
    const float* bias; T* out; int N;
    __device__ void operator()(int m, int n, float acc, int, int) const { stf(out, (size_t)m * N + n, acc + bias[n]); }
};
// dx = base (+ extra) + acc   or   dx += acc
template <typename T>
struct EpDgrad {
    const T* base; const T* extra; T* out; int N; int accumulate;
    __device__ void operator()(int m, int n, float acc, int, int) const {
        const size_t i = (size_t)m * N + n;
        float v = accumulate ? ldf(out, i) : ldf(base, i) + (extra ? ldf(extra, i) : 0.f);
        stf(out, i, v + acc);
    }
};
// wgrad operands for tap z: A(m=co, k=row) = dz[row, co]; B(k=row, n=ci) = x[row shifted by (z-half)*dil, ci]
template <typename T>
struct WgradA {
    static constexpr bool kContig = false;
    const T* dz; int C;
    __device__ float operator()(int m, int k, int) const { return ldf(dz, (size_t)k * C + m); }
};
template <typename T>
struct WgradB {
    static constexpr bool kContig = false;
    const T* x; int L, C, dil, half;
    __device__ float operator()(int k, int n, int z) const {
        const int b = k / L, l = k - b * L;
        const int l2 = l + (z - half) * dil;
        return (l2 >= 0 && l2 < L) ? ldf(x, ((size_t)b * L + l2) * C + n) : 0.f;
    }
};
struct EpWgradAtomic {
    float* dW; int C, ks;
    __device__ void operator()(int m, int n, float acc, int, int z) const {
        atomicAdd(&dW[((size_t)m * C + n) * ks + z], acc);
    }
};
struct EpAtomic {
    float* out; long ld;
    __device__ void operator()(int m, int n, float acc, int, int) const { atomicAdd(&out[(size_t)m * ld + n], acc); }
};
// out is a stack of per-head [ntot, dv] matrices: row m = head*dv + j of the [G, C] product goes to out[head][n][j]
struct EpAtomicHeads {
    float* out; int dv, ntot;
    __device__ void operator()(int m, int n, float acc, int, int) const {
        atomicAdd(&out[(size_t)(m / dv) * ((size_t)ntot * dv) + (size_t)n * dv + (m % dv)], acc);
    }
};
template <typename T>
struct EpLocalLinear {
    const float* bias; const T* h; T* z; T* u2; int N;
    __device__ void operator()(int m, int n, float acc, int, int) const {
        const size_t i = (size_t)m * N + n;
        const float zz = acc + bias[n];
        stf(z, i, zz);
        stf(u2, i, ldf(h, i) + gelu_f(rnd(z, zz)));
    }
};
template <typename T>
struct EpAddStore {
    const T* base; T* out; int N;
    __device__ void operator()(int m, int n, float acc, int, int) const {
        const size_t i = (size_t)m * N + n;
        stf(out, i, ldf(base, i) + acc);
    }
};
struct EpPool {
    float* pool; int L, G;
    __device__ void operator()(int m, int n, float acc, int, int) const {
        atomicAdd(&pool[(size_t)(m / L) * G + n], gelu_f(acc));
    }
};
template <typename T>
struct EpPoolBwd {
    const float* dpool; T* dzv; int L, G;
    __device__ void operator()(int m, int n, float acc, int, int) const {
        stf(dzv, (size_t)m * G + n, dpool[(size_t)(m / L) * G + n] * gelu_grad_f(acc));
    }
};

template <class AL, class BL, class EP>
void gemm_auto(AL al, BL bl, EP ep, int M, int N, int K, int nsplit, int zb, cudaStream_t st) {
    if (N >= 128 && M >= 128) launch_gemm_simt<128, 128, 8, 8, 8>(al, bl, ep, M, N, K, nsplit, zb, st);
    else launch_gemm_simt<64, 64, 16, 4, 4>(al, bl, ep, M, N, K, nsplit, zb, st);
    pb_count_launch(1);
}

// ---------------------------------------------------------------------------------------------
// elementwise / reduction kernels.  grid = (chunks, B): a CTA never straddles two samples.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void conv_fuse_kernel(const T* __restrict__ x, const T* __restrict__ zn, const T* __restrict__ zw,
                                 const float* __restrict__ g2l, T* __restrict__ u, double* __restrict__ stats, int L, int C) {
    __shared__ double red[32];
    const int b = blockIdx.y;
    const long long n = (long long)L * C;
    const long long per = (n + gridDim.x - 1) / gridDim.x;
    const long long beg = blockIdx.x * per, end = min(n, beg + per);
    float s = 0.f, q = 0.f;
    for (long long i = beg + threadIdx.x; i < end; i += blockDim.x) {
        const size_t gi = (size_t)b * n + i;
        const int c = (int)(i % C);
        float v = ldf(x, gi) + gelu_f(ldf(zn, gi));        // summed left to right as modules.py:211
        v += gelu_f(ldf(zw, gi));
        v += g2l[(size_t)b * C + c];
        v = rnd(u, v);
        stf(u, gi, v);
        s += v;
        q += v * v;
    }
    const double ds = block_sum((double)s, red), dq = block_sum((double)q, red);
    if (threadIdx.x == 0) { atomicAdd(&stats[2 * b], ds); atomicAdd(&stats[2 * b + 1], dq); }
}

template <typename T>
__global__ void ln_stats_kernel(const T* __restrict__ u, double* __restrict__ stats, int L, int C) {
    __shared__ double red[32];
    const int b = blockIdx.y;
    const long long n = (long long)L * C;
    const long long per = (n + gridDim.x - 1) / gridDim.x;
    const long long beg = blockIdx.x * per, end = min(n, beg + per);
    float s = 0.f, q = 0.f;
    for (long long i = beg + threadIdx.x; i < end; i += blockDim.x) {
        const float v = ldf(u, (size_t)b * n + i);
        s += v;
        q += v * v;
    }
    const double ds = block_sum((double)s, red), dq = block_sum((double)q, red);
    if (threadIdx.x == 0) { atomicAdd(&stats[2 * b], ds); atomicAdd(&stats[2 * b + 1], dq); }
}

template <typename T>
__global__ void ln_apply_kernel(const T* __restrict__ u, const double* __restrict__ stats, const float* __restrict__ gamma,
                                const float* __restrict__ beta, T* __restrict__ h, int L, int C, float eps) {
    const int b = blockIdx.y;
    const long long n = (long long)L * C;
    const LnStat st = ln_stat(stats, b, (double)n, eps);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const size_t gi = (size_t)b * n + i;
        stf(h, gi, (ldf(u, gi) - st.mean) * st.rstd * gamma[i] + beta[i]);
    }
}

// pass 1 of the joint-LN backward: each thread owns (l,c) positions and walks the batch, so dgamma/dbeta
// need no atomics; the per-sample sums s1 = sum dy*gamma, s2 = sum dy*gamma*xhat go through shared memory.
template <typename T>
__global__ void ln_bwd_pass1_kernel(const T* __restrict__ dy, const T* __restrict__ u, const double* __restrict__ stats,
                                    const float* __restrict__ gamma, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                    double* __restrict__ sums, int B, int L, int C, float eps) {
    extern __shared__ float sm[];            // [B][2] partial sums + [B][2] mean/rstd
    float* part = sm;
    float* mr = sm + 2 * B;
    const long long n = (long long)L * C;
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        const LnStat st = ln_stat(stats, b, (double)n, eps);
        mr[2 * b] = st.mean; mr[2 * b + 1] = st.rstd;
        part[2 * b] = 0.f; part[2 * b + 1] = 0.f;
    }
    __syncthreads();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < n;
    const float g = live ? gamma[i] : 0.f;
    float dg = 0.f, db = 0.f;
    for (int b = 0; b < B; ++b) {
        float a1 = 0.f, a2 = 0.f;
        if (live) {
            const size_t gi = (size_t)b * n + i;
            const float d = ldf(dy, gi);
            const float xh = (ldf(u, gi) - mr[2 * b]) * mr[2 * b + 1];
            dg += d * xh;
            db += d;
            a1 = d * g;
            a2 = a1 * xh;
        }
        a1 = warp_sum(a1);
        a2 = warp_sum(a2);
        if ((threadIdx.x & 31) == 0) { atomicAdd(&part[2 * b], a1); atomicAdd(&part[2 * b + 1], a2); }
    }
    if (live) { dgamma[i] = dg; dbeta[i] = db; }
    __syncthreads();
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        atomicAdd(&sums[2 * b], (double)part[2 * b]);
        atomicAdd(&sums[2 * b + 1], (double)part[2 * b + 1]);
    }
}

template <typename T>
__global__ void ln_bwd_pass2_kernel(const T* __restrict__ dy, const T* __restrict__ u, const double* __restrict__ stats,
                                    const float* __restrict__ gamma, const double* __restrict__ sums, T* __restrict__ du,
                                    int L, int C, float eps) {
    const int b = blockIdx.y;
    const long long n = (long long)L * C;
    const LnStat st = ln_stat(stats, b, (double)n, eps);
    const float m1 = (float)(sums[2 * b] / (double)n), m2 = (float)(sums[2 * b + 1] / (double)n);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const size_t gi = (size_t)b * n + i;
        const float xh = (ldf(u, gi) - st.mean) * st.rstd;
        stf(du, gi, st.rstd * (ldf(dy, gi) * gamma[i] - m1 - xh * m2));
    }
}

// dz = dy * gelu'(z) for up to two (z, dz) pairs; column sums of dy per sample (-> dsample[b,c]) and of
// each dz over all rows (-> dbias[c]).  block = (32, 8): x walks channels, y walks rows; grid = (chunks, B).
template <typename T>
__global__ void gelu_bwd_colsum_kernel(const T* __restrict__ dy, const T* __restrict__ z0, const T* __restrict__ z1,
                                       T* __restrict__ dz0, T* __restrict__ dz1, float* __restrict__ dsample,
                                       float* __restrict__ db0, float* __restrict__ db1, int L, int C) {
    extern __shared__ float sm[];            // [3][C]
    const int b = blockIdx.y;
    const int rows_per = (L + gridDim.x - 1) / gridDim.x;
    const int l0 = blockIdx.x * rows_per, l1 = min(L, l0 + rows_per);
    for (int i = threadIdx.y * 32 + threadIdx.x; i < 3 * C; i += 256) sm[i] = 0.f;
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += 32) {
        float sy = 0.f, s0 = 0.f, s1 = 0.f;
        for (int l = l0 + threadIdx.y; l < l1; l += 8) {
            const size_t gi = ((size_t)b * L + l) * C + c;
            const float d = ldf(dy, gi);
            sy += d;
            const float g0 = d * gelu_grad_f(ldf(z0, gi));
            stf(dz0, gi, g0);
            s0 += rnd(dz0, g0);
            if (z1) {
                const float g1 = d * gelu_grad_f(ldf(z1, gi));
                stf(dz1, gi, g1);
                s1 += rnd(dz1, g1);
            }
        }
        atomicAdd(&sm[c], sy);
        atomicAdd(&sm[C + c], s0);
        atomicAdd(&sm[2 * C + c], s1);
    }
    __syncthreads();
    for (int c = threadIdx.y * 32 + threadIdx.x; c < C; c += 256) {
        if (dsample) atomicAdd(&dsample[(size_t)b * C + c], sm[c]);
        atomicAdd(&db0[c], sm[C + c]);
        if (z1) atomicAdd(&db1[c], sm[2 * C + c]);
    }
}


// ---------------------------------------------------------------------------------------------
// vectorised (8 elements per thread) versions of the HBM-bound kernels, used whenever C % 8 == 0
// ---------------------------------------------------------------------------------------------
template <typename T> struct V8;
template <> struct V8<float> {
    static __device__ __forceinline__ void ld(const float* p, float (&v)[8]) {
        const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
    static __device__ __forceinline__ void st(float* p, const float (&v)[8]) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    static __device__ __forceinline__ float gelu_grad(float z) { return gelu_grad_f(z); }
    static __device__ __forceinline__ void gelu_grad8(float (&z)[8]) {          // in place
#pragma unroll
        for (int j = 0; j < 8; ++j) z[j] = gelu_grad_f(z[j]);
    }
    struct Raw { float4 a, b; };
    static __device__ __forceinline__ Raw ld_raw(const float* p) { return Raw{*reinterpret_cast<const float4*>(p), *reinterpret_cast<const float4*>(p + 4)}; }
    static __device__ __forceinline__ void unpack(const Raw& r, float (&v)[8]) {
        v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
    }
};
template <> struct V8<bf16> {
    static __device__ __forceinline__ void ld(const bf16* p, float (&v)[8]) {
        const uint4 r = *reinterpret_cast<const uint4*>(p);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
    }
    // stores the rounded values and returns them in v (sums are taken of stored values)
    static __device__ __forceinline__ void st(bf16* p, float (&v)[8]) {
        uint4 r;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y;
        }
        *reinterpret_cast<uint4*>(p) = r;
    }
    static __device__ __forceinline__ float gelu_grad(float z) { return gelu_grad_fast(z); }
    static __device__ __forceinline__ void gelu_grad8(float (&z)[8]) { gelu_grad_fast_n(z); }      // in place, packed pairs
    typedef uint4 Raw;
    static __device__ __forceinline__ Raw ld_raw(const bf16* p) { return *reinterpret_cast<const uint4*>(p); }
    static __device__ __forceinline__ void unpack(const Raw& r, float (&v)[8]) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
    }
};
__device__ __forceinline__ void ldf8(const float* p, float (&v)[8]) { V8<float>::ld(p, v); }

// per-sample RAW sums s1 = sum dy*gamma, t2 = sum dy*gamma*u.  grid = (chunks, B).  The consumer turns t2 into
// sum dy*gamma*xhat = rstd*(t2 - mean*s1) (ln_m2_from_raw), so this pass needs neither the statistics nor any double
// arithmetic before its loads -- it is one round trip to memory plus a reduction.
template <typename T>
__global__ void ln_bwd_sums_vec_kernel(const T* __restrict__ dy, const T* __restrict__ u, const float* __restrict__ gamma,
                                       double* __restrict__ sums, long long n) {
    __shared__ double red[32];
    pdl_wait();
    const int b = blockIdx.y;
    const long long nv = n / 8;
    float s1 = 0.f, s2 = 0.f;
    const long long step = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + step < nv; i += 2 * step) {
        float d0[8], x0[8], g0[8], d1[8], x1[8], g1[8];
        V8<T>::ld(dy + (size_t)b * n + i * 8, d0);
        V8<T>::ld(u + (size_t)b * n + i * 8, x0);
        V8<T>::ld(dy + (size_t)b * n + (i + step) * 8, d1);
        V8<T>::ld(u + (size_t)b * n + (i + step) * 8, x1);
        ldf8(gamma + i * 8, g0);
        ldf8(gamma + (i + step) * 8, g1);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float a = d0[j] * g0[j], c = d1[j] * g1[j];
            s1 += a + c;
            s2 += a * x0[j] + c * x1[j];
        }
    }
    for (; i < nv; i += step) {
        float d[8], x[8], g[8];
        V8<T>::ld(dy + (size_t)b * n + i * 8, d);
        V8<T>::ld(u + (size_t)b * n + i * 8, x);
        ldf8(gamma + i * 8, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float a = d[j] * g[j];
            s1 += a;
            s2 += a * x[j];
        }
    }
    const double d1 = block_sum((double)s1, red), d2 = block_sum((double)s2, red);
    if (threadIdx.x == 0) { atomicAdd(&sums[2 * b], d1); atomicAdd(&sums[2 * b + 1], d2); }
}
// (sum dy*gamma*xhat) / n from the raw sums, in double: xhat = (u - mean) * rstd
__device__ __forceinline__ float ln_m2_from_raw(const double* stats, int b, double n, float eps, double s1, double t2) {
    const double mean = stats[2 * b] / n;
    double var = stats[2 * b + 1] / n - mean * mean;
    if (var < 0.0) var = 0.0;
    return (float)((t2 - mean * s1) * rsqrt(var + (double)eps) / n);
}

// du = rstd*(dy*gamma - m1 - xhat*m2); dgamma/dbeta accumulated over this CTA's group of samples.
// grid = (ceil(nv/256), groups); thread owns one 8-wide (l,c) vector and walks `per_group` samples.
template <typename T>
__global__ void ln_bwd_apply_vec_kernel(const T* __restrict__ dy, const T* __restrict__ u, const double* __restrict__ stats,
                                        const float* __restrict__ gamma, const double* __restrict__ sums, T* __restrict__ du,
                                        float* __restrict__ dgamma, float* __restrict__ dbeta, int B, long long n, float eps,
                                        int per_group, int use_atomics) {
    extern __shared__ float prm[];       // [per_group][4]: mean, rstd, m1, m2
    pdl_wait();
    const int b0 = blockIdx.y * per_group;
    const int nb = min(per_group, B - b0);
    for (int k = threadIdx.x; k < nb; k += blockDim.x) {
        const LnStat st = ln_stat(stats, b0 + k, (double)n, eps);
        prm[4 * k] = st.mean; prm[4 * k + 1] = st.rstd;
        prm[4 * k + 2] = (float)(sums[2 * (b0 + k)] / (double)n);
        prm[4 * k + 3] = ln_m2_from_raw(stats, b0 + k, (double)n, eps, sums[2 * (b0 + k)], sums[2 * (b0 + k) + 1]);
    }
    __syncthreads();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n / 8) return;
    float g[8], ag[8], ab[8];
    ldf8(gamma + i * 8, g);
#pragma unroll
    for (int j = 0; j < 8; ++j) { ag[j] = 0.f; ab[j] = 0.f; }
    for (int k = 0; k < nb; ++k) {
        const size_t off = (size_t)(b0 + k) * n + i * 8;
        float d[8], x[8], o[8];
        V8<T>::ld(dy + off, d);
        V8<T>::ld(u + off, x);
        const float mean = prm[4 * k], rstd = prm[4 * k + 1], m1 = prm[4 * k + 2], m2 = prm[4 * k + 3];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float xh = (x[j] - mean) * rstd;
            o[j] = rstd * (d[j] * g[j] - m1 - xh * m2);
            ag[j] += d[j] * xh;
            ab[j] += d[j];
        }
        V8<T>::st(du + off, o);
    }
    if (use_atomics) {       // 16-byte vector reductions: 4 instead of 16 atomics per thread
        atomicAdd(reinterpret_cast<float4*>(dgamma + i * 8), make_float4(ag[0], ag[1], ag[2], ag[3]));
        atomicAdd(reinterpret_cast<float4*>(dgamma + i * 8 + 4), make_float4(ag[4], ag[5], ag[6], ag[7]));
        atomicAdd(reinterpret_cast<float4*>(dbeta + i * 8), make_float4(ab[0], ab[1], ab[2], ab[3]));
        atomicAdd(reinterpret_cast<float4*>(dbeta + i * 8 + 4), make_float4(ab[4], ab[5], ab[6], ab[7]));
    } else {
        V8<float>::st(dgamma + i * 8, ag);
        V8<float>::st(dbeta + i * 8, ab);
    }
}

// dz = dy * gelu'(z) for one or two z tensors, plus per-sample column sums of dy and global column sums of dz.
// block = 256 threads = (C/8) channel vectors x rows; grid = (row chunks, B).
template <typename T>
__global__ void gelu_bwd_colsum_vec_kernel(const T* __restrict__ dy, const T* __restrict__ z0, const T* __restrict__ z1,
                                           T* __restrict__ dz0, T* __restrict__ dz1, float* __restrict__ dsample,
                                           float* __restrict__ db0, float* __restrict__ db1, int L, int C) {
    extern __shared__ float sm[];        // [3][C]
    pdl_wait();
    const int b = blockIdx.y;
    const int vpr = C / 8;                        // vectors per row
    const int rows_pb = blockDim.x / vpr;         // rows handled concurrently
    const int tv = threadIdx.x % vpr, tr = threadIdx.x / vpr;
    const int rows_per = (L + gridDim.x - 1) / gridDim.x;
    const int l0 = blockIdx.x * rows_per, l1 = min(L, l0 + rows_per);
    for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sm[i] = 0.f;
    __syncthreads();
    float sy[8], s0[8], s1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { sy[j] = 0.f; s0[j] = 0.f; s1[j] = 0.f; }
    if (tr < rows_pb) {
        for (int l = l0 + tr; l < l1; l += rows_pb) {
            const size_t off = ((size_t)b * L + l) * C + tv * 8;
            float d[8], z[8], o[8];
            V8<T>::ld(dy + off, d);
            V8<T>::ld(z0 + off, z);
            V8<T>::gelu_grad8(z);
#pragma unroll
            for (int j = 0; j < 8; ++j) { sy[j] += d[j]; o[j] = d[j] * z[j]; }
            V8<T>::st(dz0 + off, o);
#pragma unroll
            for (int j = 0; j < 8; ++j) s0[j] += o[j];
            if (z1) {
                V8<T>::ld(z1 + off, z);
                V8<T>::gelu_grad8(z);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = d[j] * z[j];
                V8<T>::st(dz1 + off, o);
#pragma unroll
                for (int j = 0; j < 8; ++j) s1[j] += o[j];
            }
        }
        // threads that share a channel vector sit vpr lanes apart: fold the rows held inside a warp with shuffles,
        // then one shared-memory add per warp and vector
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            for (int o = 16; o >= vpr && o >= 1; o >>= 1) {
                sy[j] += __shfl_xor_sync(0xffffffffu, sy[j], o);
                s0[j] += __shfl_xor_sync(0xffffffffu, s0[j], o);
                s1[j] += __shfl_xor_sync(0xffffffffu, s1[j], o);
            }
        }
        if ((threadIdx.x & 31) < vpr || vpr >= 32) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(&sm[tv * 8 + j], sy[j]);
                atomicAdd(&sm[C + tv * 8 + j], s0[j]);
                if (z1) atomicAdd(&sm[2 * C + tv * 8 + j], s1[j]);
            }
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        if (dsample) atomicAdd(&dsample[(size_t)b * C + c], sm[c]);
        atomicAdd(&db0[c], sm[C + c]);
        if (z1) atomicAdd(&db1[c], sm[2 * C + c]);
    }
}

// LayerNorm input gradient and the GELU gradients that follow it, in one pass over the activations:
//   du = rstd*(dy*gamma - m1 - xhat*m2)   dz0 = du*gelu'(z0)   [dz1 = du*gelu'(z1)]
// plus dgamma/dbeta (per (l,c), over the batch), db0/db1 (per channel, over batch and positions) and, if asked,
// dsample[b,c] = sum_l du.  Same thread layout as ln_bwd_apply_vec_kernel: a thread owns one 8-wide (l,c) vector and
// walks its group's samples, so the (l,c) and per-channel sums stay in registers; per-sample sums go through one
// shared-memory slot per warp and sample.  Requires (C/8) to divide 32.
template <typename T, bool TWO>
__global__ void __launch_bounds__(256)
ln_bwd_apply_gelu_kernel(const T* __restrict__ dy, const T* __restrict__ u, const double* __restrict__ stats,
                         const float* __restrict__ gamma, const double* __restrict__ sums, const T* __restrict__ z0,
                         const T* __restrict__ z1, T* __restrict__ du, T* __restrict__ dz0, T* __restrict__ dz1,
                         float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dsample,
                         float* __restrict__ db0, float* __restrict__ db1, int B, int C, long long n, float eps, int per_group) {
    extern __shared__ float sh[];        // [per_group][4] LN parameters | [8 warps][per_group + 2][C] partial column sums
    float* prm = sh;
    float* part = sh + 4 * per_group;
    pdl_wait();
    const int b0 = blockIdx.y * per_group;
    const int nb = min(per_group, B - b0);
    for (int k = threadIdx.x; k < nb; k += blockDim.x) {
        const LnStat st = ln_stat(stats, b0 + k, (double)n, eps);
        prm[4 * k] = st.mean; prm[4 * k + 1] = st.rstd;
        prm[4 * k + 2] = (float)(sums[2 * (b0 + k)] / (double)n);
        prm[4 * k + 3] = ln_m2_from_raw(stats, b0 + k, (double)n, eps, sums[2 * (b0 + k)], sums[2 * (b0 + k) + 1]);
    }
    __syncthreads();
    const int vpr = C / 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slots = per_group + 2;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n / 8;
    const int tv = (int)(i % vpr);
    float* wpart = part + (size_t)warp * slots * C;
    float g[8], ag[8], ab[8], c0[8], c1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { g[j] = 0.f; ag[j] = 0.f; ab[j] = 0.f; c0[j] = 0.f; c1[j] = 0.f; }
    if (valid) ldf8(gamma + i * 8, g);
    // the next sample's operands are requested before this sample's arithmetic: twice the bytes in flight per thread
    typename V8<T>::Raw rd, rx, ra, rb;
    if (valid && nb > 0) {
        const size_t off = (size_t)b0 * n + i * 8;
        rd = V8<T>::ld_raw(dy + off); rx = V8<T>::ld_raw(u + off); ra = V8<T>::ld_raw(z0 + off);
        if (TWO) rb = V8<T>::ld_raw(z1 + off);
    }
    for (int k = 0; k < nb; ++k) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = 0.f;
        if (valid) {
            const size_t off = (size_t)(b0 + k) * n + i * 8;
            float d[8], x[8], za[8], zb[8], e[8];
            V8<T>::unpack(rd, d); V8<T>::unpack(rx, x); V8<T>::unpack(ra, za);
            if (TWO) V8<T>::unpack(rb, zb);
            if (k + 1 < nb) {
                rd = V8<T>::ld_raw(dy + off + n); rx = V8<T>::ld_raw(u + off + n); ra = V8<T>::ld_raw(z0 + off + n);
                if (TWO) rb = V8<T>::ld_raw(z1 + off + n);
            }
            const float mean = prm[4 * k], rstd = prm[4 * k + 1], m1 = prm[4 * k + 2], m2 = prm[4 * k + 3];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float xh = (x[j] - mean) * rstd;
                o[j] = rstd * (d[j] * g[j] - m1 - xh * m2);
                ag[j] += d[j] * xh;
                ab[j] += d[j];
            }
            V8<T>::st(du + off, o);                  // o now holds the stored (rounded) values
            V8<T>::gelu_grad8(za);
#pragma unroll
            for (int j = 0; j < 8; ++j) e[j] = o[j] * za[j];
            V8<T>::st(dz0 + off, e);
#pragma unroll
            for (int j = 0; j < 8; ++j) c0[j] += e[j];
            if (TWO) {
                V8<T>::gelu_grad8(zb);
#pragma unroll
                for (int j = 0; j < 8; ++j) e[j] = o[j] * zb[j];
                V8<T>::st(dz1 + off, e);
#pragma unroll
                for (int j = 0; j < 8; ++j) c1[j] += e[j];
            }
        }
        if (dsample != nullptr) {        // fold the rows this warp holds, one slot per warp and sample
#pragma unroll
            for (int j = 0; j < 8; ++j)
                for (int s = 16; s >= vpr; s >>= 1) o[j] += __shfl_xor_sync(0xffffffffu, o[j], s);
            if (lane < vpr) {
#pragma unroll
                for (int j = 0; j < 8; ++j) wpart[(size_t)(2 + k) * C + tv * 8 + j] = o[j];
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
        for (int s = 16; s >= vpr; s >>= 1) {
            c0[j] += __shfl_xor_sync(0xffffffffu, c0[j], s);
            if (TWO) c1[j] += __shfl_xor_sync(0xffffffffu, c1[j], s);
        }
    if (lane < vpr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            wpart[tv * 8 + j] = c0[j];
            wpart[C + tv * 8 + j] = c1[j];
        }
    }
    if (valid) {
        atomicAdd(reinterpret_cast<float4*>(dgamma + i * 8), make_float4(ag[0], ag[1], ag[2], ag[3]));
        atomicAdd(reinterpret_cast<float4*>(dgamma + i * 8 + 4), make_float4(ag[4], ag[5], ag[6], ag[7]));
        atomicAdd(reinterpret_cast<float4*>(dbeta + i * 8), make_float4(ab[0], ab[1], ab[2], ab[3]));
        atomicAdd(reinterpret_cast<float4*>(dbeta + i * 8 + 4), make_float4(ab[4], ab[5], ab[6], ab[7]));
    }
    __syncthreads();
    // across the 8 warps: slot 0 -> db0, slot 1 -> db1, slot 2+k -> dsample[b0+k]
    const int first = 0, last = dsample != nullptr ? 2 + nb : 2;
    for (int e = threadIdx.x; e < (last - first) * C; e += blockDim.x) {
        const int slot = e / C, c = e % C;
        if (slot == 1 && !TWO) continue;
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += part[((size_t)w * slots + slot) * C + c];
        float* dst = slot == 0 ? db0 + c : slot == 1 ? db1 + c : dsample + (size_t)(b0 + slot - 2) * C + c;
        atomicAdd(dst, s);
    }
}

template <typename T>
int launch_gelu_bwd_colsum(const void* dy, const void* z0, const void* z1, void* dz0, void* dz1, float* dsample, float* db0,
                           float* db1, int B, int L, int C, cudaStream_t st) {
    if (C % 8 == 0 && C / 8 <= 256 && 256 % (C / 8) == 0) {
        dim3 grid(max(1, min(ceil_div(L, 64), 16)), B);
        PB_CUDA(pb_launch_pdl(gelu_bwd_colsum_vec_kernel<T>, grid, dim3(256), 3 * C * sizeof(float), st, (const T*)dy, (const T*)z0,
                              (const T*)z1, (T*)dz0, (T*)dz1, dsample, db0, db1, L, C));
    } else {
        dim3 grid(max(1, min(ceil_div(L, 64), 16)), B);
        gelu_bwd_colsum_kernel<T><<<grid, dim3(32, 8), 3 * C * sizeof(float), st>>>((const T*)dy, (const T*)z0, (const T*)z1, (T*)dz0,
                                                                                   (T*)dz1, dsample, db0, db1, L, C);
    }
    pb_count_launch(1);
    return PB_OK;
}

int split_for(int tiles, long K) {
    int s = 592 / (tiles > 0 ? tiles : 1);
    if (s < 1) s = 1;
    while (s > 1 && K / s < 256) --s;
    return s;
}

}  // namespace

extern "C" {

int pb_ln_joint_stats(const void* u, double* stats, int B, int L, int C, int dt, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0, PB_ERR_SHAPE, "ln_joint_stats: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    PB_CUDA(cudaMemsetAsync(stats, 0, (size_t)B * 2 * sizeof(double), st));
    dim3 grid(max(1, min(ceil_div((long)L * C, 4096), 16)), B);
    DISPATCH_DT(dt, (ln_stats_kernel<T><<<grid, 256, 0, st>>>((const T*)u, stats, L, C)));
    pb_count_launch(1);
    return pb_check_launch("pb_ln_joint_stats");
}

int pb_ln_joint_apply(const void* u, const double* stats, const float* gamma, const float* beta, void* h, int B, int L, int C,
                      float eps, int dt, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0, PB_ERR_SHAPE, "ln_joint_apply: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    if (dt == PB_BF16 && pb_tc_ln_apply_ok(L, C)) return pb_tc_ln_apply(u, stats, gamma, beta, h, B, L, C, eps, st);
    dim3 grid(max(1, min(ceil_div((long)L * C, 1024), 32)), B);
    DISPATCH_DT(dt, (ln_apply_kernel<T><<<grid, 256, 0, st>>>((const T*)u, stats, gamma, beta, (T*)h, L, C, eps)));
    pb_count_launch(1);
    return pb_check_launch("pb_ln_joint_apply");
}

int pb_ln_joint_bwd(const void* dy, const void* u, const double* stats, const float* gamma, void* du, float* dgamma,
                    float* dbeta, double* sums, int B, int L, int C, float eps, int dt, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0, PB_ERR_SHAPE, "ln_joint_bwd: bad shape");
    PB_REQUIRE(B <= 4096, PB_ERR_SHAPE, "ln_joint_bwd: batch %d exceeds the 4096 supported per call", B);
    cudaStream_t st = (cudaStream_t)stream;
    const long long n = (long long)L * C;
    if (n % 8 == 0) {
        const long long nv = n / 8;
        // enough sample groups to put >= ~150k threads in flight
        int groups = (int)((75776 + nv - 1) / nv);
        if (groups > B) groups = B;
        if (groups < 1) groups = 1;
        const int per_group = (B + groups - 1) / groups;
        groups = (B + per_group - 1) / per_group;
        dim3 g1((unsigned)max(1, min(ceil_div(nv, 256), 16)), B);
        dim3 g2((unsigned)ceil_div(nv, 256), groups);
        DISPATCH_DT(dt, {
            PB_CUDA(pb_launch_pdl(ln_bwd_sums_vec_kernel<T>, g1, dim3(256), 0, st, (const T*)dy, (const T*)u, gamma, sums, n));
            PB_CUDA(pb_launch_pdl(ln_bwd_apply_vec_kernel<T>, g2, dim3(256), 4 * per_group * sizeof(float), st, (const T*)dy, (const T*)u,
                                  stats, gamma, (const double*)sums, (T*)du, dgamma, dbeta, B, n, eps, per_group, groups > 1 ? 1 : 0));
        });
    } else {
        DISPATCH_DT(dt, (ln_bwd_pass1_kernel<T><<<ceil_div(n, 128), 128, 4 * B * sizeof(float), st>>>(
                            (const T*)dy, (const T*)u, stats, gamma, dgamma, dbeta, sums, B, L, C, eps)));
        dim3 grid(max(1, min(ceil_div(n, 1024), 32)), B);
        DISPATCH_DT(dt, (ln_bwd_pass2_kernel<T><<<grid, 256, 0, st>>>((const T*)dy, (const T*)u, stats, gamma, sums, (T*)du, L, C, eps)));
    }
    pb_count_launch(2);
    return pb_check_launch("pb_ln_joint_bwd");
}

int pb_ln_joint_bwd_gelu_ok(int B, int L, int C) {
    const int vpr = C / 8;
    return B > 0 && L > 0 && C > 0 && C % 8 == 0 && vpr <= 32 && 32 % vpr == 0 && B <= 4096;
}

int pb_ln_joint_bwd_gelu(const void* dy, const void* u, const double* stats, const float* gamma, const void* z0, const void* z1,
                         void* du, void* dz0, void* dz1, float* dgamma, float* dbeta, float* dsample, float* db0, float* db1,
                         double* sums, int B, int L, int C, float eps, int dt, void* stream) {
    PB_REQUIRE(pb_ln_joint_bwd_gelu_ok(B, L, C), PB_ERR_SHAPE, "ln_joint_bwd_gelu: unsupported shape B=%d L=%d C=%d (see pb_ln_joint_bwd_gelu_ok)", B, L, C);
    PB_REQUIRE(z0 != nullptr && dz0 != nullptr && db0 != nullptr, PB_ERR_SHAPE, "ln_joint_bwd_gelu: z0, dz0 and db0 are required");
    PB_REQUIRE((z1 == nullptr) == (dz1 == nullptr) && (z1 == nullptr) == (db1 == nullptr), PB_ERR_SHAPE,
               "ln_joint_bwd_gelu: z1, dz1 and db1 go together");
    cudaStream_t st = (cudaStream_t)stream;
    const long long n = (long long)L * C;
    const long long nv = n / 8;
    // 128 registers x 256 threads: two CTAs per SM, 296 resident CTAs; stay within ONE wave (a second, partial wave of
    // CTAs that each walk their samples serially would double the time)
    int groups = (int)(296 / ((nv + 255) / 256));
    if (groups > B) groups = B;
    if (groups < 1) groups = 1;
    int per_group = (B + groups - 1) / groups;
    const int cap = max(1, (12288 - 16 * C) / (8 * C + 4));      // keeps the partial-sum slots inside 48 KB of shared memory
    if (per_group > cap) per_group = cap;
    groups = (B + per_group - 1) / per_group;
    const size_t smem = (size_t)(4 * per_group + 8 * (per_group + 2) * C) * sizeof(float);
    dim3 g1((unsigned)max(1, min(ceil_div(nv, 256), 16)), B);
    dim3 g2((unsigned)ceil_div(nv, 256), groups);
    DISPATCH_DT(dt, {
        PB_CUDA(pb_launch_pdl(ln_bwd_sums_vec_kernel<T>, g1, dim3(256), 0, st, (const T*)dy, (const T*)u, gamma, sums, n));
        if (z1 != nullptr) {
            PB_CUDA(pb_launch_pdl(ln_bwd_apply_gelu_kernel<T, true>, g2, dim3(256), smem, st, (const T*)dy, (const T*)u, stats, gamma,
                                  (const double*)sums, (const T*)z0, (const T*)z1, (T*)du, (T*)dz0, (T*)dz1, dgamma, dbeta, dsample, db0,
                                  db1, B, C, n, eps, per_group));
        } else {
            PB_CUDA(pb_launch_pdl(ln_bwd_apply_gelu_kernel<T, false>, g2, dim3(256), smem, st, (const T*)dy, (const T*)u, stats, gamma,
                                  (const double*)sums, (const T*)z0, (const T*)z1, (T*)du, (T*)dz0, (T*)dz1, dgamma, dbeta, dsample, db0,
                                  db1, B, C, n, eps, per_group));
        }
    });
    pb_count_launch(2);
    return pb_check_launch("pb_ln_joint_bwd_gelu");
}

int pb_conv_pair_fwd(const void* x, const void* wn, const void* ww, const float* bn, const float* bw, const float* g2l,
                     void* zn, void* zw, void* u, double* stats, int B, int L, int C, int ksize, int dil, int dt,
                     void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0 && ksize > 0 && (ksize & 1) && dil > 0, PB_ERR_SHAPE,
               "conv_pair_fwd: bad shape B=%d L=%d C=%d k=%d dil=%d (odd kernel sizes only: 'same' padding)", B, L, C,
               ksize, dil);
    cudaStream_t st = (cudaStream_t)stream;
    if (dt == PB_BF16 && pb_tc_conv_ok(B, L, C, ksize, dil))
        return pb_tc_conv_pair_fwd(x, wn, ww, bn, bw, g2l, zn, zw, u, stats, B, L, C, ksize, dil, st);
    const int M = B * L, K = ksize * C, half = ksize / 2;
    DISPATCH_DT(dt, {
        gemm_auto(ConvA<T>{(const T*)x, L, C, 1, half, 1}, ConvPackB<T>{(const T*)wn, C}, EpBiasStore<T>{bn, (T*)zn, C}, M, C, K, 1, 1, st);
        gemm_auto(ConvA<T>{(const T*)x, L, C, dil, half, 1}, ConvPackB<T>{(const T*)ww, C}, EpBiasStore<T>{bw, (T*)zw, C}, M, C, K, 1, 1, st);
        dim3 grid(max(1, min(ceil_div((long)L * C, 4096), 16)), B);
        conv_fuse_kernel<T><<<grid, 256, 0, st>>>((const T*)x, (const T*)zn, (const T*)zw, g2l, (T*)u, stats, L, C);
        pb_count_launch(1);
    });
    return pb_check_launch("pb_conv_pair_fwd");
}

int pb_conv_pair_bwd(const void* du, const void* zn, const void* zw, const void* x, const void* wn_bwd, const void* ww_bwd,
                     void* dzn, void* dzw, void* dx, float* dWn, float* dWw, float* dbn, float* dbw, float* dg2l, int B,
                     int L, int C, int ksize, int dil, int dt, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0 && ksize > 0 && (ksize & 1) && dil > 0, PB_ERR_SHAPE, "conv_pair_bwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    const int M = B * L, K = ksize * C, half = ksize / 2;
    const bool tc = dt == PB_BF16 && pb_tc_conv_ok(B, L, C, ksize, dil);
    // dx == NULL skips the data-gradient half, dWn == NULL the weight-gradient half (the caller may run them on two streams)
    const bool do_dx = dx != nullptr, do_dw = dWn != nullptr;
    PB_REQUIRE(do_dx || do_dw, PB_ERR_SHAPE, "conv_pair_bwd: nothing to do (dx and dWn both NULL)");
    PB_REQUIRE(!do_dw || dWw != nullptr, PB_ERR_SHAPE, "conv_pair_bwd: dWn and dWw go together");
    if (!tc && do_dw) {
        PB_CUDA(cudaMemsetAsync(dWn, 0, (size_t)C * C * ksize * sizeof(float), st));
        PB_CUDA(cudaMemsetAsync(dWw, 0, (size_t)C * C * ksize * sizeof(float), st));
    }
    DISPATCH_DT(dt, {
        if (zn != nullptr) {
            PB_REQUIRE(do_dx, PB_ERR_SHAPE, "conv_pair_bwd: the GELU gradients belong to the data-gradient call (pass zn = NULL with dx = NULL)");
            launch_gelu_bwd_colsum<T>(du, zn, zw, dzn, dzw, dg2l, dbn, dbw, B, L, C, st);
        }
        if (tc) {
            int rc = do_dx ? pb_tc_conv_pair_dgrad(du, dzn, dzw, wn_bwd, ww_bwd, dx, B, L, C, ksize, dil, st) : 0;
            if (rc) return rc;
            rc = do_dw ? pb_tc_conv_wgrad(dzn, dzw, x, dWn, dWw, B, L, C, ksize, dil, ws, ws_bytes, st) : 0;
            if (rc) return rc;
        } else {
            if (do_dx) {
                // data gradient: same implicit GEMM with the taps mirrored (sign = -1) and the [t][ci][co] pack
                gemm_auto(ConvA<T>{(const T*)dzn, L, C, 1, half, -1}, ConvPackB<T>{(const T*)wn_bwd, C},
                          EpDgrad<T>{(const T*)du, nullptr, (T*)dx, C, 0}, M, C, K, 1, 1, st);
                gemm_auto(ConvA<T>{(const T*)dzw, L, C, dil, half, -1}, ConvPackB<T>{(const T*)ww_bwd, C},
                          EpDgrad<T>{nullptr, nullptr, (T*)dx, C, 1}, M, C, K, 1, 1, st);
            }
            if (do_dw) {
                // weight gradient: one [C x C] GEMM per tap, contraction over all B*L rows, split-K with atomics
                const int tiles = ceil_div(C, C >= 128 ? 128 : 64) * ceil_div(C, C >= 128 ? 128 : 64) * ksize;
                const int ns = split_for(tiles, M);
                gemm_auto(WgradA<T>{(const T*)dzn, C}, WgradB<T>{(const T*)x, L, C, 1, half}, EpWgradAtomic{dWn, C, ksize}, C, C, M, ns, ksize, st);
                gemm_auto(WgradA<T>{(const T*)dzw, C}, WgradB<T>{(const T*)x, L, C, dil, half}, EpWgradAtomic{dWw, C, ksize}, C, C, M, ns, ksize, st);
            }
        }
    });
    return pb_check_launch("pb_conv_pair_bwd");
}

int pb_local_linear_fwd(const void* h, const void* W, const float* bias, void* z, void* u2, double* stats2, int B, int L, int C,
                        int dt, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0, PB_ERR_SHAPE, "local_linear_fwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    if (dt == PB_BF16 && pb_tc_linear_ok(B, L, C))
        return pb_tc_local_linear_fwd(h, W, bias, z, u2, stats2, B, L, C, st);
    const int M = B * L;
    DISPATCH_DT(dt, {
        gemm_auto(RowMajorA<T>{(const T*)h, C}, WeightNK<T>{(const T*)W, C}, EpLocalLinear<T>{bias, (const T*)h, (T*)z, (T*)u2, C}, M, C, C, 1, 1, st);
        dim3 grid(max(1, min(ceil_div((long)L * C, 4096), 16)), B);
        ln_stats_kernel<T><<<grid, 256, 0, st>>>((const T*)u2, stats2, L, C);
        pb_count_launch(1);
    });
    return pb_check_launch("pb_local_linear_fwd");
}

int pb_local_linear_bwd(const void* du2, const void* z, const void* h, const void* W, const void* Wt, void* dz, void* dh, float* dW, float* db,
                        int B, int L, int C, int dt, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0, PB_ERR_SHAPE, "local_linear_bwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    const int M = B * L;
    const bool tc = dt == PB_BF16 && Wt != nullptr && pb_tc_linear_ok(B, L, C);
    // dh == NULL skips the data-gradient half, dW == NULL the weight-gradient half (the caller may run them on two streams)
    const bool do_dx = dh != nullptr, do_dw = dW != nullptr;
    PB_REQUIRE(do_dx || do_dw, PB_ERR_SHAPE, "local_linear_bwd: nothing to do (dh and dW both NULL)");
    if (!tc && do_dw) PB_CUDA(cudaMemsetAsync(dW, 0, (size_t)C * C * sizeof(float), st));
    DISPATCH_DT(dt, {
        if (z != nullptr) {
            PB_REQUIRE(do_dx, PB_ERR_SHAPE, "local_linear_bwd: the GELU gradient belongs to the data-gradient call (pass z = NULL with dh = NULL)");
            launch_gelu_bwd_colsum<T>(du2, z, nullptr, dz, nullptr, nullptr, db, nullptr, B, L, C, st);
        }
        if (tc) {
            int rc = pb_tc_local_linear_bwd(du2, dz, h, Wt, dh, dW, B, L, C, ws, ws_bytes, st);
            if (rc) return rc;
        } else {
            // dh = du2 + dz W
            if (do_dx) gemm_auto(RowMajorA<T>{(const T*)dz, C}, MatKN<T>{(const T*)W, C}, EpAddStore<T>{(const T*)du2, (T*)dh, C}, M, C, C, 1, 1, st);
            // dW[n,c] = sum_rows dz[row,n] h[row,c]
            const int tiles = ceil_div(C, C >= 128 ? 128 : 64) * ceil_div(C, C >= 128 ? 128 : 64);
            if (do_dw) gemm_auto(ColMajorA<T>{(const T*)dz, C}, MatKN<T>{(const T*)h, C}, EpAtomic{dW, C}, C, C, M, split_for(tiles, M), 1, st);
        }
    });
    return pb_check_launch("pb_local_linear_bwd");
}

int pb_attn_pool_fwd(const void* h, const void* wvT, float* pool, int B, int L, int C, int G, int dt, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0 && G > 0, PB_ERR_SHAPE, "attn_pool_fwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    if (dt == PB_BF16 && pb_tc_attn_ok(B, L, C, G)) return pb_tc_attn_pool_fwd(h, wvT, pool, B, L, C, G, st);
    DISPATCH_DT(dt, gemm_auto(RowMajorA<T>{(const T*)h, C}, WeightNK<T>{(const T*)wvT, C}, EpPool{pool, L, G}, B * L, G, C, 1, 1, st));
    return pb_check_launch("pb_attn_pool_fwd");
}

int pb_attn_pool_bwd(const void* h, const void* wvT, const void* wv, const float* dpool, const void* dh_in, void* dzv, void* dh_out,
                     float* dWvT, int head_dv, int B, int L, int C, int G, int dt, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0 && G > 0, PB_ERR_SHAPE, "attn_pool_bwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    const int M = B * L;
    if (dt == PB_BF16 && wv != nullptr && pb_tc_attn_ok(B, L, C, G))
        return pb_tc_attn_pool_bwd(h, wvT, wv, dpool, dh_in, dzv, dh_out, dWvT, head_dv, B, L, C, G, ws, ws_bytes, st);
    if (dWvT) PB_CUDA(cudaMemsetAsync(dWvT, 0, (size_t)G * C * sizeof(float), st));
    DISPATCH_DT(dt, {
        if (dh_out != nullptr) {       // dh_out == NULL: weight-gradient half only, dzv is then an input
            gemm_auto(RowMajorA<T>{(const T*)h, C}, WeightNK<T>{(const T*)wvT, C}, EpPoolBwd<T>{dpool, (T*)dzv, L, G}, M, G, C, 1, 1, st);
            gemm_auto(RowMajorA<T>{(const T*)dzv, G}, MatKN<T>{(const T*)wvT, C}, EpAddStore<T>{(const T*)dh_in, (T*)dh_out, C}, M, C, G, 1, 1, st);
        }
        if (dWvT) {
            const int tiles = ceil_div(G, 128) * ceil_div(C, 128);
            if (head_dv > 0)
                gemm_auto(ColMajorA<T>{(const T*)dzv, G}, MatKN<T>{(const T*)h, C}, EpAtomicHeads{dWvT, head_dv, C}, G, C, M, split_for(tiles, M), 1, st);
            else
                gemm_auto(ColMajorA<T>{(const T*)dzv, G}, MatKN<T>{(const T*)h, C}, EpAtomic{dWvT, C}, G, C, M, split_for(tiles, M), 1, st);
        }
    });
    return pb_check_launch("pb_attn_pool_bwd");
}

int pb_tc_path(const char* op, int B, int L, int C, int G, int ksize, int dil) {
    if (op == nullptr) return 0;
    if (!strcmp(op, "conv")) {
        if (!pb_tc_conv_ok(B, L, C, ksize, dil)) return 0;
        return (pb_tc_mode() != 4 && pb_tc_conv2_ok(B, L, C, ksize, dil)) ? 2 : 1;
    }
    if (!strcmp(op, "local_linear")) return pb_tc_linear_ok(B, L, C) ? 1 : 0;
    if (!strcmp(op, "attn_pool")) {
        if (!pb_tc_attn_ok(B, L, C, G)) return 0;
        return pb_tc_pool_ok(B, L, C, G) ? 2 : 1;
    }
    return 0;
}

}  // extern "C"
