// GlobalAttention as the reference EXECUTES it (modules.py:21-92), one fused launch per block: per head
//     q  = tanh(g Wq)                      the K query rows are copies of one vector (x_global is repeated, :51), so one is formed
//     kt = tanh(h Wk)            [L, K]
//     s  = q . kt^T / sqrt(K)    [L]       identical for each of the K query rows
//     P  = softmax(s)                      axis 0: over the repeated-query axis, the reference's nn.Softmax(dim=1) (:34,:58)
//                                          axis 1: over the sequence positions, the paper's reading (softmax_axis="sequence")
//     v  = gelu(h Wv)            [L, Dv]
//     out = P v                  [Dv]      (every query row gives the same vector)
//   att[b, head*Dv + d] = sum(W_parameter) * out[d]                          (:87-92 with K identical rows)
// Scores, probabilities and values live in shared memory / registers; only q [B,H,K], P [B,H,L] and the pooled vector
// [B,G] are kept for backward, which recomputes kt and v tile by tile.
//
// Under axis 0 every probability is exp(s - s) / (K exp(0)) = 1/K exactly, the layer collapses to (sum(Wp)/K) sum_l gelu(h Wv)
// and Wq / Wk receive exactly zero gradient: that closed form is what pb_attn_pool_* evaluate on the tensor cores (the
// default fast path).  This file is the op-faithful kernel: it checks the closed form (tests) and is the only
// implementation of axis 1, where Wq and Wk matter.  fp32 arithmetic, SIMT; grid = (batch, heads).
#include "common.cuh"
#include "../../include/pb_b200.h"

namespace {

constexpr int kT = 256;
constexpr int kTL = 16;            // positions per tile

template <typename T>
__device__ __forceinline__ void load_tile(float* h_s, const T* __restrict__ h, int l0, int L, int C) {
    for (int i = threadIdx.x; i < kTL * C; i += kT) {
        const int l = i / C, c = i - l * C;
        h_s[i] = (l0 + l < L) ? ldf(h, (size_t)(l0 + l) * C + c) : 0.f;
    }
}

__device__ __forceinline__ float block_reduce_sum(float v, float* red) { return block_sum(v, red); }
__device__ __forceinline__ float block_reduce_max(float v, float* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float r = red[0];
    for (int i = 1; i < kT / 32; ++i) r = fmaxf(r, red[i]);
    __syncthreads();
    return r;
}

struct AttnShape { int B, L, C, G, K, H, Dv, axis; };

// shared memory: q[K] | s[L] (scores, then probabilities) | h tile [kTL][C] | red[32] | products [kTL][K]
template <typename T>
__global__ void __launch_bounds__(kT) attention_fwd_kernel(const T* __restrict__ h, const float* __restrict__ g, const float* __restrict__ Wq,
                                                           const float* __restrict__ Wk, const float* __restrict__ Wv,
                                                           const float* __restrict__ Wp, float* __restrict__ att, float* __restrict__ q_out,
                                                           float* __restrict__ p_out, float* __restrict__ pool_out, AttnShape sh) {
    extern __shared__ __align__(16) float sm[];
    const int b = blockIdx.x, hd = blockIdx.y;
    const int L = sh.L, C = sh.C, G = sh.G, K = sh.K, Dv = sh.Dv;
    float* q_s = sm;
    float* s_s = q_s + K;
    float* h_s = s_s + L;
    float* red = h_s + kTL * C;
    float* scr = red + 32;                                   // [kTL][K]
    const T* hb = h + (size_t)b * L * C;
    const float* wq = Wq + (size_t)hd * G * K;
    const float* wk = Wk + (size_t)hd * C * K;
    const float* wv = Wv + (size_t)hd * C * Dv;
    // ---- q = tanh(g Wq)  (:51-53)
    for (int k = threadIdx.x; k < K; k += kT) {
        float a = 0.f;
        for (int j = 0; j < G; ++j) a = fmaf(g[(size_t)b * G + j], wq[(size_t)j * K + k], a);
        const float qv = tanhf(a);
        q_s[k] = qv;
        q_out[((size_t)b * sh.H + hd) * K + k] = qv;
    }
    __syncthreads();
    const float inv_sqrt_k = rsqrtf((float)K);
    // ---- scores s[l] = q . tanh(h[l] Wk) / sqrt(K)  (:54,:57-58)
    for (int l0 = 0; l0 < L; l0 += kTL) {
        __syncthreads();
        load_tile(h_s, hb, l0, L, C);
        __syncthreads();
        for (int i = threadIdx.x; i < kTL * K; i += kT) {    // (position, key) pairs
            const int l = i / K, k = i - l * K;
            float a = 0.f;
            for (int c = 0; c < C; ++c) a = fmaf(h_s[l * C + c], __ldg(wk + (size_t)c * K + k), a);
            scr[i] = q_s[k] * tanhf(a);
        }
        __syncthreads();
        for (int l = threadIdx.x; l < kTL; l += kT) {        // fold the K products of each position
            if (l0 + l < L) {
                float tot = 0.f;
                for (int k = 0; k < K; ++k) tot += scr[l * K + k];
                s_s[l0 + l] = tot * inv_sqrt_k;
            }
        }
    }
    __syncthreads();
    // ---- probabilities
    if (sh.axis == 1) {                                      // softmax over the sequence positions
        float m = -INFINITY;
        for (int l = threadIdx.x; l < L; l += kT) m = fmaxf(m, s_s[l]);
        m = block_reduce_max(m, red);
        float z = 0.f;
        for (int l = threadIdx.x; l < L; l += kT) { const float e = __expf(s_s[l] - m); s_s[l] = e; z += e; }
        z = block_reduce_sum(z, red);
        const float iz = 1.0f / z;
        for (int l = threadIdx.x; l < L; l += kT) s_s[l] *= iz;
    } else {                                                 // softmax over the K identical query rows (:34): exp(s - s) / (K exp(0))
        for (int l = threadIdx.x; l < L; l += kT) {
            const float e = __expf(s_s[l] - s_s[l]);
            s_s[l] = e / ((float)K * e);
        }
    }
    __syncthreads();
    for (int l = threadIdx.x; l < L; l += kT) p_out[((size_t)b * sh.H + hd) * L + l] = s_s[l];
    // ---- out[d] = sum_l P[l] gelu(h[l] Wv[:, d])  (:55,:59-60) and the head mix (:92)
    float wsum = 0.f;
    for (int k = threadIdx.x; k < K; k += kT) wsum += Wp[k];
    wsum = block_reduce_sum(wsum, red);
    for (int d0 = 0; d0 < Dv; d0 += kT) {
        const int d = d0 + threadIdx.x;
        float acc = 0.f;
        for (int l0 = 0; l0 < L; l0 += kTL) {
            __syncthreads();
            load_tile(h_s, hb, l0, L, C);
            __syncthreads();
            if (d < Dv) {
                for (int l = 0; l < kTL && l0 + l < L; ++l) {
                    float z = 0.f;
                    for (int c = 0; c < C; ++c) z = fmaf(h_s[l * C + c], __ldg(wv + (size_t)c * Dv + d), z);
                    acc = fmaf(s_s[l0 + l], gelu_f(z), acc);
                }
            }
        }
        if (d < Dv) {
            pool_out[(size_t)b * G + hd * Dv + d] = acc;
            att[(size_t)b * G + hd * Dv + d] = wsum * acc;
        }
    }
}

// ---- backward ------------------------------------------------------------------------------------------------------------
// shared memory: q[K] | p[L] | dp[L] (then ds) | h tile [kTL][C] | dz tile [kTL][DC] (also dzk [kTL][K]) | dW accumulators
// [C][DC] (also [C][K]) | red[32].  DC = columns of the value matrix handled per pass (<= 128).
template <typename T>
__global__ void __launch_bounds__(kT) attention_bwd_kernel(const T* __restrict__ h, const float* __restrict__ g, const float* __restrict__ Wq,
                                                           const float* __restrict__ Wk, const float* __restrict__ Wv,
                                                           const float* __restrict__ Wp, const float* __restrict__ datt,
                                                           const float* __restrict__ q_in, const float* __restrict__ p_in,
                                                           const float* __restrict__ pool_in, float* __restrict__ dh, float* __restrict__ dg,
                                                           float* __restrict__ dWq, float* __restrict__ dWk, float* __restrict__ dWv,
                                                           float* __restrict__ dWp_scalar, AttnShape sh, int DC) {
    extern __shared__ __align__(16) float sm[];
    const int b = blockIdx.x, hd = blockIdx.y;
    const int L = sh.L, C = sh.C, G = sh.G, K = sh.K, Dv = sh.Dv;
    const int WA = DC > K ? DC : K;
    float* q_s = sm;
    float* p_s = q_s + K;
    float* dp_s = p_s + L;
    float* h_s = dp_s + L;
    float* dz_s = h_s + kTL * C;                             // [kTL][WA]
    float* dW_s = dz_s + kTL * WA;                           // [C][WA]
    float* red = dW_s + C * WA;
    const T* hb = h + (size_t)b * L * C;
    const float* wq = Wq + (size_t)hd * G * K;
    const float* wk = Wk + (size_t)hd * C * K;
    const float* wv = Wv + (size_t)hd * C * Dv;
    float* dhb = dh + (size_t)b * L * C;
    for (int k = threadIdx.x; k < K; k += kT) q_s[k] = q_in[((size_t)b * sh.H + hd) * K + k];
    for (int l = threadIdx.x; l < L; l += kT) { p_s[l] = p_in[((size_t)b * sh.H + hd) * L + l]; dp_s[l] = 0.f; }
    float wsum = 0.f;
    for (int k = threadIdx.x; k < K; k += kT) wsum += Wp[k];
    wsum = block_reduce_sum(wsum, red);
    // d sum(Wp): every entry of W_parameter gets sum_{b,g} datt * pooled
    {
        float t = 0.f;
        for (int d = threadIdx.x; d < Dv; d += kT) t += datt[(size_t)b * G + hd * Dv + d] * pool_in[(size_t)b * G + hd * Dv + d];
        t = block_reduce_sum(t, red);
        if (threadIdx.x == 0 && t != 0.f) atomicAdd(dWp_scalar, t);
    }
    // ---- pass A: the value path, DC columns at a time
    for (int d0 = 0; d0 < Dv; d0 += DC) {
        const int dc = min(DC, Dv - d0);
        for (int e = threadIdx.x; e < C * dc; e += kT) dW_s[e] = 0.f;
        for (int l0 = 0; l0 < L; l0 += kTL) {
            __syncthreads();
            load_tile(h_s, hb, l0, L, C);
            __syncthreads();
            // z, v, dp, dz for (position, column) pairs
            for (int i = threadIdx.x; i < kTL * dc; i += kT) {
                const int l = i / dc, d = i - l * dc;
                float dzv = 0.f;
                if (l0 + l < L) {
                    float z = 0.f;
                    for (int c = 0; c < C; ++c) z = fmaf(h_s[l * C + c], __ldg(wv + (size_t)c * Dv + d0 + d), z);
                    const float dout = wsum * datt[(size_t)b * G + hd * Dv + d0 + d];
                    atomicAdd(&dp_s[l0 + l], dout * gelu_f(z));
                    dzv = dout * p_s[l0 + l] * gelu_grad_f(z);
                }
                dz_s[l * WA + d] = dzv;
            }
            __syncthreads();
            // dWv[c, d] += sum_l h[l, c] dz[l, d]
            for (int e = threadIdx.x; e < C * dc; e += kT) {
                const int c = e / dc, d = e - c * dc;
                float a = dW_s[e];
                for (int l = 0; l < kTL; ++l) a = fmaf(h_s[l * C + c], dz_s[l * WA + d], a);
                dW_s[e] = a;
            }
            // dh[l, c] += sum_d dz[l, d] Wv[c, d]
            for (int i = threadIdx.x; i < kTL * C; i += kT) {
                const int l = i / C, c = i - l * C;
                if (l0 + l < L) {
                    float a = 0.f;
                    for (int d = 0; d < dc; ++d) a = fmaf(dz_s[l * WA + d], __ldg(wv + (size_t)c * Dv + d0 + d), a);
                    atomicAdd(&dhb[(size_t)(l0 + l) * C + c], a);
                }
            }
        }
        __syncthreads();
        for (int e = threadIdx.x; e < C * dc; e += kT) {
            const int c = e / dc, d = e - c * dc;
            atomicAdd(&dWv[((size_t)hd * C + c) * Dv + d0 + d], dW_s[e]);
        }
        __syncthreads();
    }
    if (sh.axis != 1) return;       // softmax over identical rows: d scores = 0, so Wq, Wk and the layer's dg receive exactly zero
    // ---- softmax backward over the sequence: ds = P (dp - sum P dp)
    float t = 0.f;
    for (int l = threadIdx.x; l < L; l += kT) t += p_s[l] * dp_s[l];
    t = block_reduce_sum(t, red);
    for (int l = threadIdx.x; l < L; l += kT) dp_s[l] = p_s[l] * (dp_s[l] - t);
    __syncthreads();
    // ---- pass B: the key path.  s = q . kt / sqrt(K): dq[k] = sum_l ds kt / sqrt(K); dkt = ds q / sqrt(K); dzk = dkt (1 - kt^2)
    const float inv_sqrt_k = rsqrtf((float)K);
    float* dq_s = red + 32;                                  // [K]
    for (int k = threadIdx.x; k < K; k += kT) dq_s[k] = 0.f;
    for (int e = threadIdx.x; e < C * K; e += kT) dW_s[e] = 0.f;
    for (int l0 = 0; l0 < L; l0 += kTL) {
        __syncthreads();
        load_tile(h_s, hb, l0, L, C);
        __syncthreads();
        for (int i = threadIdx.x; i < kTL * K; i += kT) {
            const int l = i / K, k = i - l * K;
            float dzk = 0.f;
            if (l0 + l < L) {
                float a = 0.f;
                for (int c = 0; c < C; ++c) a = fmaf(h_s[l * C + c], __ldg(wk + (size_t)c * K + k), a);
                const float kt = tanhf(a), ds = dp_s[l0 + l];
                atomicAdd(&dq_s[k], ds * kt * inv_sqrt_k);
                dzk = ds * q_s[k] * inv_sqrt_k * (1.f - kt * kt);
            }
            dz_s[l * WA + k] = dzk;
        }
        __syncthreads();
        for (int e = threadIdx.x; e < C * K; e += kT) {
            const int c = e / K, k = e - c * K;
            float a = dW_s[e];
            for (int l = 0; l < kTL; ++l) a = fmaf(h_s[l * C + c], dz_s[l * WA + k], a);
            dW_s[e] = a;
        }
        for (int i = threadIdx.x; i < kTL * C; i += kT) {
            const int l = i / C, c = i - l * C;
            if (l0 + l < L) {
                float a = 0.f;
                for (int k = 0; k < K; ++k) a = fmaf(dz_s[l * WA + k], __ldg(wk + (size_t)c * K + k), a);
                atomicAdd(&dhb[(size_t)(l0 + l) * C + c], a);
            }
        }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < C * K; e += kT) atomicAdd(&dWk[(size_t)hd * C * K + e], dW_s[e]);
    // ---- query path: q = tanh(g Wq): dzq = dq (1 - q^2); dWq[j, k] += g[j] dzq[k]; dg[j] += sum_k dzq[k] Wq[j, k]
    for (int k = threadIdx.x; k < K; k += kT) dq_s[k] = dq_s[k] * (1.f - q_s[k] * q_s[k]);
    __syncthreads();
    for (int e = threadIdx.x; e < G * K; e += kT) {
        const int j = e / K, k = e - j * K;
        atomicAdd(&dWq[(size_t)hd * G * K + e], g[(size_t)b * G + j] * dq_s[k]);
    }
    for (int j = threadIdx.x; j < G; j += kT) {
        float a = 0.f;
        for (int k = 0; k < K; ++k) a = fmaf(dq_s[k], wq[(size_t)j * K + k], a);
        atomicAdd(&dg[(size_t)b * G + j], a);
    }
}

__global__ void fill_wp_kernel(float* dWp, const float* scalar, int K) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < K) dWp[k] = *scalar;
}

}  // namespace

extern "C" {

size_t pb_attention_smem(int L, int C, int K, int Dv, int backward) {
    if (!backward) return ((size_t)K + L + (size_t)kTL * C + 32 + (size_t)kTL * K) * sizeof(float);
    const int DC = Dv < 128 ? Dv : 128, WA = DC > K ? DC : K;
    return ((size_t)K + 2 * (size_t)L + (size_t)kTL * C + (size_t)kTL * WA + (size_t)C * WA + 32 + K) * sizeof(float);
}

/* Wq [H, G, K], Wk [H, C, K], Wv [H, C, Dv] are the heads' matrices stacked (modules.py:36-47); axis 0 = the reference's
 * softmax axis, 1 = sequence positions.  q_out [B,H,K], p_out [B,H,L], pool_out [B,G] are kept for backward. */
int pb_attention_fwd(const void* h_act, const float* g, const float* Wq, const float* Wk, const float* Wv, const float* Wp, float* att,
                     float* q_out, float* p_out, float* pool_out, int B, int L, int C, int G, int K, int H, int axis, int dt,
                     void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0 && G > 0 && K > 0 && H > 0 && G % H == 0 && (axis == 0 || axis == 1), PB_ERR_SHAPE,
               "attention_fwd: bad shape B=%d L=%d C=%d G=%d K=%d H=%d axis=%d", B, L, C, G, K, H, axis);
    AttnShape sh{B, L, C, G, K, H, G / H, axis};
    const size_t smem = pb_attention_smem(L, C, K, G / H, 0);
    PB_REQUIRE(smem <= 200 * 1024, PB_ERR_UNSUPPORTED, "attention_fwd: %zu bytes of shared memory needed", smem);
    cudaStream_t st = (cudaStream_t)stream;
    if (dt == PB_BF16) {
        PB_CUDA(cudaFuncSetAttribute(attention_fwd_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attention_fwd_kernel<bf16><<<dim3(B, H), kT, smem, st>>>((const bf16*)h_act, g, Wq, Wk, Wv, Wp, att, q_out, p_out, pool_out, sh);
    } else if (dt == PB_F32) {
        PB_CUDA(cudaFuncSetAttribute(attention_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attention_fwd_kernel<float><<<dim3(B, H), kT, smem, st>>>((const float*)h_act, g, Wq, Wk, Wv, Wp, att, q_out, p_out, pool_out, sh);
    } else {
        pb_set_error("attention_fwd: unknown dtype code %d", dt);
        return PB_ERR_UNSUPPORTED;
    }
    pb_count_launch(1);
    return pb_check_launch("pb_attention_fwd");
}

/* dh [B,L,C] float, dg [B,G], dWq, dWk, dWv (stacked like the weights) are ACCUMULATED into: zero on entry.  dWp [K] is
 * written.  scratch: one float, zero on entry.  Under axis 0, dg / dWq / dWk stay zero (their true gradient is 0). */
int pb_attention_bwd(const void* h_act, const float* g, const float* Wq, const float* Wk, const float* Wv, const float* Wp,
                     const float* datt, const float* q, const float* p, const float* pool, float* dh, float* dg, float* dWq, float* dWk,
                     float* dWv, float* dWp, float* scratch, int B, int L, int C, int G, int K, int H, int axis, int dt, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0 && G > 0 && K > 0 && H > 0 && G % H == 0 && (axis == 0 || axis == 1), PB_ERR_SHAPE,
               "attention_bwd: bad shape");
    AttnShape sh{B, L, C, G, K, H, G / H, axis};
    const int Dv = G / H, DC = Dv < 128 ? Dv : 128;
    const size_t smem = pb_attention_smem(L, C, K, Dv, 1);
    PB_REQUIRE(smem <= 200 * 1024, PB_ERR_UNSUPPORTED, "attention_bwd: %zu bytes of shared memory needed", smem);
    cudaStream_t st = (cudaStream_t)stream;
    if (dt == PB_BF16) {
        PB_CUDA(cudaFuncSetAttribute(attention_bwd_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attention_bwd_kernel<bf16><<<dim3(B, H), kT, smem, st>>>((const bf16*)h_act, g, Wq, Wk, Wv, Wp, datt, q, p, pool, dh, dg, dWq, dWk,
                                                                  dWv, scratch, sh, DC);
    } else if (dt == PB_F32) {
        PB_CUDA(cudaFuncSetAttribute(attention_bwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attention_bwd_kernel<float><<<dim3(B, H), kT, smem, st>>>((const float*)h_act, g, Wq, Wk, Wv, Wp, datt, q, p, pool, dh, dg, dWq, dWk,
                                                                   dWv, scratch, sh, DC);
    } else {
        pb_set_error("attention_bwd: unknown dtype code %d", dt);
        return PB_ERR_UNSUPPORTED;
    }
    fill_wp_kernel<<<(K + 127) / 128, 128, 0, st>>>(dWp, scratch, K);
    pb_count_launch(2);
    return pb_check_launch("pb_attention_bwd");
}

}  // extern "C"
