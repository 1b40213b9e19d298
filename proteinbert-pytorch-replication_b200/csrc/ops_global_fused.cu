// Fused global-track kernels: the whole global side of one ProteinBERTBlock in FOUR launches
// (forward: pre, post; backward: post, pre) instead of ~37.  All fp32 SIMT -- the global track is < 1 % of
// the FLOPs but is cancellation-prone, and at M = batch rows it is pure launch latency.
//
//   pre  (needs only the block input g):   zg = g W1^T + b1, g2l = gelu(zg)      (modules.py:208)
//                                          z3 = g W3^T + b3, gh  = gelu(z3)      (modules.py:221)
//   post (needs the attention pooling):    g1 = LN(g + gh + coef*pool)           (modules.py:92,223-224)
//                                          z4 = g1 W4^T + b4
//                                          g2 = LN(g1 + gelu(z4))                (modules.py:226-229)
//
// Structure of every kernel: each CTA owns a slice of 8 output columns.  Whatever is row-wise and needed by
// all slices (a LayerNorm of all rows, an activation derivative) is recomputed by every CTA into shared
// memory, where it becomes the A operand [rows x K] of the CTA's small GEMM; what needs complete rows of a
// GEMM result (the second LayerNorm, the first LayerNorm's backward) is done by the LAST CTA to finish
// (threadfence + atomic ticket), so no grid-wide barrier or cooperative launch is needed.
#include "common.cuh"
#include "../../include/pb_b200.h"

namespace {

constexpr int kNT = 8;            // output columns per CTA
constexpr int kThreads = 256;
constexpr float kEps = 1e-5f;

__device__ __forceinline__ int row_tile_of(int K) { return K <= 704 ? 64 : (K <= 1472 ? 32 : 16); }
static int host_row_tile(int K) { return K <= 704 ? 64 : (K <= 1472 ? 32 : 16); }

// ---- CTA-level product: out[r][j] = sum_k A_s[r*lda + k] * Bt_s[k*8 + j],  r < MT, j < 8, k < Kd ---------
// thread (r = tid % MT, ks = tid / MT) accumulates the k = ks, ks + nks, ... slice; partials are combined through
// red_s [nks][MT][8].  The result is returned in out[] for threads with ks == 0 (tid < MT).
__device__ __forceinline__ void cta_gemm8(const float* A_s, int lda, const float* Bt_s, int Kd, int MT, float* red_s,
                                          float (&out)[kNT]) {
    const int r = threadIdx.x % MT, ks = threadIdx.x / MT, nks = kThreads / MT;
    float acc[kNT];
#pragma unroll
    for (int j = 0; j < kNT; ++j) acc[j] = 0.f;
    const float* a = A_s + (size_t)r * lda;
    for (int k = ks; k < Kd; k += nks) {
        const float av = a[k];
        const float4 b0 = *reinterpret_cast<const float4*>(Bt_s + k * kNT);
        const float4 b1 = *reinterpret_cast<const float4*>(Bt_s + k * kNT + 4);
        acc[0] = fmaf(av, b0.x, acc[0]); acc[1] = fmaf(av, b0.y, acc[1]);
        acc[2] = fmaf(av, b0.z, acc[2]); acc[3] = fmaf(av, b0.w, acc[3]);
        acc[4] = fmaf(av, b1.x, acc[4]); acc[5] = fmaf(av, b1.y, acc[5]);
        acc[6] = fmaf(av, b1.z, acc[6]); acc[7] = fmaf(av, b1.w, acc[7]);
    }
#pragma unroll
    for (int j = 0; j < kNT; ++j) red_s[((size_t)ks * MT + r) * kNT + j] = acc[j];
    __syncthreads();
    if (ks == 0) {
#pragma unroll
        for (int j = 0; j < kNT; ++j) {
            float s = 0.f;
            for (int q = 0; q < nks; ++q) s += red_s[((size_t)q * MT + r) * kNT + j];
            out[j] = s;
        }
    }
    __syncthreads();
}

// Bt_s[k][j] = W[(n0 + j) * ldw + k]   (weight rows are the output columns: "X W^T")
__device__ __forceinline__ void load_bt_rows(float* Bt_s, const float* __restrict__ W, int ldw, int n0, int nvalid, int Kd) {
    for (int i = threadIdx.x; i < kNT * Kd; i += kThreads) {
        const int j = i / Kd, k = i - j * Kd;
        Bt_s[k * kNT + j] = j < nvalid ? W[(size_t)(n0 + j) * ldw + k] : 0.f;
    }
}
// Bt_s[k][j] = W[k * ldw + c0 + j]     (weight columns are the output columns: "dZ W")
__device__ __forceinline__ void load_bt_cols(float* Bt_s, const float* __restrict__ W, int ldw, int c0, int nvalid, int Kd) {
    for (int i = threadIdx.x; i < kNT * Kd; i += kThreads) {
        const int k = i / kNT, j = i - k * kNT;
        Bt_s[i] = j < nvalid ? W[(size_t)k * ldw + c0 + j] : 0.f;
    }
}

// accumulate dW rows of this CTA's slice: acc[q][j] += sum_m D_s[m*8 + j] * X[(row0+m)*ldx + kcol], kcol = tid + q*256.
// Rows are taken eight at a time with all global loads issued before the FMAs (latency is L2, ~600 cycles).
template <int QMAX>
__device__ __forceinline__ void dw_accumulate(float (&acc)[QMAX][kNT], const float* D_s, const float* __restrict__ X, int ldx,
                                              int row0, int mt, int Kout) {
    for (int m0 = 0; m0 < mt; m0 += 8) {
        float x[8][QMAX];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int q = 0; q < QMAX; ++q) {
                const int kc = threadIdx.x + q * kThreads;
                x[i][q] = (m0 + i < mt && kc < Kout) ? __ldg(X + (size_t)(row0 + m0 + i) * ldx + kc) : 0.f;
            }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (m0 + i < mt) {
                const float4 d0 = *reinterpret_cast<const float4*>(D_s + (m0 + i) * kNT);
                const float4 d1 = *reinterpret_cast<const float4*>(D_s + (m0 + i) * kNT + 4);
#pragma unroll
                for (int q = 0; q < QMAX; ++q) {
                    const float xv = x[i][q];
                    acc[q][0] = fmaf(d0.x, xv, acc[q][0]); acc[q][1] = fmaf(d0.y, xv, acc[q][1]);
                    acc[q][2] = fmaf(d0.z, xv, acc[q][2]); acc[q][3] = fmaf(d0.w, xv, acc[q][3]);
                    acc[q][4] = fmaf(d1.x, xv, acc[q][4]); acc[q][5] = fmaf(d1.y, xv, acc[q][5]);
                    acc[q][6] = fmaf(d1.z, xv, acc[q][6]); acc[q][7] = fmaf(d1.w, xv, acc[q][7]);
                }
            }
        }
    }
}

// Cooperative fill of a [MT x K] shared tile (row pitch lda) from row-major global tensors, eight float4 per thread
// in flight: F(row, k, i) returns the float4 of source i at (row, k..k+3); G(vals) combines them into the float4 to
// store.  Rows >= mt are zero-filled.  K % 4 == 0.
template <int NSRC, class Load, class Combine>
__device__ __forceinline__ void fill_tile(float* A_s, int lda, int MT, int mt, int K, Load load, Combine combine) {
    const int kv = K / 4, total = MT * kv;
    constexpr int U = 4;
    for (int base = threadIdx.x; base < total; base += kThreads * U) {
        float4 v[U][NSRC];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * kThreads;
            const int r = i / kv, k = (i - r * kv) * 4;
            if (i < total && r < mt) {
#pragma unroll
                for (int sidx = 0; sidx < NSRC; ++sidx) v[u][sidx] = load(r, k, sidx);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * kThreads;
            if (i < total) {
                const int r = i / kv, k = (i - r * kv) * 4;
                float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < mt) o = combine(v[u], r, k);
                float* d = A_s + (size_t)r * lda + k;
                d[0] = o.x; d[1] = o.y; d[2] = o.z; d[3] = o.w;
            }
        }
    }
}
__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 ld4cg(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

template <int QMAX>
__device__ __forceinline__ void dw_store(const float (&acc)[QMAX][kNT], float* __restrict__ dW, int ldw, int n0, int nvalid, int Kout) {
#pragma unroll
    for (int q = 0; q < QMAX; ++q) {
        const int kc = threadIdx.x + q * kThreads;
        if (kc < Kout)
#pragma unroll
            for (int j = 0; j < kNT; ++j)
                if (j < nvalid) dW[(size_t)(n0 + j) * ldw + kc] = acc[q][j];
    }
}
constexpr int kQ = 4;             // supports contraction-output widths up to 1024 per dW row

__device__ __forceinline__ float coef_of(const float* __restrict__ Wp, int Kdim, float* red /* >= 32 */) {
    float t = 0.f;
    for (int k = threadIdx.x; k < Kdim; k += kThreads) t += Wp[k];
    return block_sum(t, red) / (float)Kdim;
}

// last-CTA ticket: returns true in every thread of the CTA that arrives last; resets the counter for reuse
__device__ __forceinline__ bool last_cta_done(int* counter, int* flag_s) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int ticket = atomicAdd(counter, 1);
        *flag_s = (ticket == (int)gridDim.x - 1);
        if (*flag_s) *counter = 0;
    }
    __syncthreads();
    const bool last = *flag_s != 0;
    if (last) __threadfence();
    return last;
}

// ============================================================================================================
// pre forward:  [zg | z3] = g [W1 ; W3]^T + [b1 ; b3];  g2l = gelu(zg), gh = gelu(z3)
// ============================================================================================================
struct PreFwd {
    const float *g, *W1, *b1, *W3, *b3;
    float *zg, *g2l, *z3, *gh;
    int B, G, C;
};

__global__ void __launch_bounds__(kThreads) global_pre_fwd_kernel(PreFwd p) {
    extern __shared__ __align__(16) float sm[];
    const int K = p.G, MT = row_tile_of(K), lda = K + 1;
    float* A_s = sm;                                   // [MT][lda]
    float* Bt_s = A_s + (size_t)MT * lda;              // [K][8]
    float* red_s = Bt_s + (size_t)K * kNT;             // [256/MT][MT][8] = [256*8]
    const int n0 = blockIdx.x * kNT;                   // column in the concatenated [C + G] output
    const bool is_g2l = n0 < p.C;                      // slices never straddle (C % 8 == 0)
    const float* W = is_g2l ? p.W1 : p.W3;
    const float* bias = is_g2l ? p.b1 : p.b3;
    const int c0 = is_g2l ? n0 : n0 - p.C;
    const int ncols = is_g2l ? p.C : p.G;
    load_bt_rows(Bt_s, W, K, c0, min(kNT, ncols - c0), K);
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();
        fill_tile<1>(A_s, lda, MT, mt, K,
                     [&](int r, int k, int) { return ld4(p.g + (size_t)(row0 + r) * K + k); },
                     [&](float4 (&v)[1], int, int) { return v[0]; });
        __syncthreads();
        float out[kNT];
        cta_gemm8(A_s, lda, Bt_s, K, MT, red_s, out);
        if (threadIdx.x < mt) {
            const int row = row0 + threadIdx.x;
            float* zo = (is_g2l ? p.zg : p.z3) + (size_t)row * ncols + c0;
            float* yo = (is_g2l ? p.g2l : p.gh) + (size_t)row * ncols + c0;
#pragma unroll
            for (int j = 0; j < kNT; ++j)
                if (c0 + j < ncols) {
                    const float z = out[j] + bias[c0 + j];
                    zo[j] = z;
                    yo[j] = gelu_f(z);
                }
        }
    }
}

// ============================================================================================================
// pre backward: dzg = dg2l*gelu'(zg), dz3 = dgh*gelu'(z3);  dg = dz3 W3 + dzg W1;  dW3 = dz3^T g, dW1 = dzg^T g
// ============================================================================================================
struct PreBwd {
    const float *dg2l, *dgh, *zg, *z3, *g, *W1, *W3;
    float *dg, *dW1, *db1, *dW3, *db3;
    int B, G, C;
};

__device__ __forceinline__ float4 mul_gelu_grad4(const float4 d, const float4 z) {
    return make_float4(d.x * gelu_grad_f(z.x), d.y * gelu_grad_f(z.y), d.z * gelu_grad_f(z.z), d.w * gelu_grad_f(z.w));
}

__global__ void __launch_bounds__(kThreads) global_pre_bwd_kernel(PreBwd p) {
    extern __shared__ __align__(16) float sm[];
    const int G = p.G, C = p.C, KA = G + C, MT = row_tile_of(KA), lda = KA + 1;
    float* A_s = sm;                                   // [MT][lda] = [dz3 | dzg]
    float* Bt_s = A_s + (size_t)MT * lda;              // [KA][8]
    float* red_s = Bt_s + (size_t)KA * kNT;            // [256*8]
    float* D_s = red_s + kThreads * kNT;               // [MT][8] slice of dz for the dW product
    // CTA i owns columns [8i, 8i+8) of dg (if < G) and rows [8i, 8i+8) of the stacked [dW3 ; dW1] matrix.
    const int n0 = blockIdx.x * kNT;                   // 0 .. G + C
    const bool has_dg = n0 < G;
    if (has_dg) {                                      // Bt = [W3[:, n0..] ; W1[:, n0..]]  (rows = contraction index)
        load_bt_cols(Bt_s, p.W3, G, n0, min(kNT, G - n0), G);
        load_bt_cols(Bt_s + (size_t)G * kNT, p.W1, G, n0, min(kNT, G - n0), C);
    }
    const bool w_is_3 = n0 < G;
    const int wrow0 = w_is_3 ? n0 : n0 - G;
    float accw[kQ][kNT];
#pragma unroll
    for (int q = 0; q < kQ; ++q)
#pragma unroll
        for (int j = 0; j < kNT; ++j) accw[q][j] = 0.f;
    float colsum = 0.f;                                // db of column (threadIdx.x < 8)
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();
        fill_tile<2>(A_s, lda, MT, mt, G,
                     [&](int r, int k, int i) { return ld4((i == 0 ? p.dgh : p.z3) + (size_t)(row0 + r) * G + k); },
                     [&](float4 (&v)[2], int, int) { return mul_gelu_grad4(v[0], v[1]); });
        fill_tile<2>(A_s + G, lda, MT, mt, C,
                     [&](int r, int k, int i) { return ld4((i == 0 ? p.dg2l : p.zg) + (size_t)(row0 + r) * C + k); },
                     [&](float4 (&v)[2], int, int) { return mul_gelu_grad4(v[0], v[1]); });
        __syncthreads();
        for (int i = threadIdx.x; i < MT * kNT; i += kThreads) {
            const int m = i / kNT, j = i - m * kNT;
            D_s[i] = A_s[(size_t)m * lda + n0 + j];    // n0 + j < KA always ((G + C) % 8 == 0)
        }
        __syncthreads();
        if (threadIdx.x < kNT)
            for (int m = 0; m < mt; ++m) colsum += D_s[m * kNT + threadIdx.x];
        dw_accumulate<kQ>(accw, D_s, p.g, G, row0, mt, G);
        if (has_dg) {
            float out[kNT];
            cta_gemm8(A_s, lda, Bt_s, KA, MT, red_s, out);
            if (threadIdx.x < mt) {
                float* o = p.dg + (size_t)(row0 + threadIdx.x) * G + n0;
#pragma unroll
                for (int j = 0; j < kNT; ++j)
                    if (n0 + j < G) o[j] = out[j];
            }
        }
    }
    dw_store<kQ>(accw, w_is_3 ? p.dW3 : p.dW1, G, wrow0, kNT, G);
    if (threadIdx.x < kNT) (w_is_3 ? p.db3 : p.db1)[wrow0 + threadIdx.x] = colsum;
}

// ============================================================================================================
// post forward
// ============================================================================================================
struct PostFwd {
    const float *g, *gh, *pool, *Wp, *gamma1, *beta1, *W4, *b4, *gamma2, *beta2;
    float *g1, *xhat1, *rstd1, *z4, *xhat2, *rstd2, *g2;
    int* counter;
    int B, G, Kdim;
};

// LayerNorm of the rows of a shared tile, one warp per row, in place: a <- xhat*gamma + beta; optional global copies
__device__ __forceinline__ void ln_rows_smem(float* A_s, int lda, int mt, int G, const float* __restrict__ gamma,
                                             const float* __restrict__ beta, float* xhat_out, float* y_out, float* rstd_out,
                                             int row0) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < mt; r += kThreads / 32) {
        float* a = A_s + (size_t)r * lda;
        float s = 0.f;
        for (int k = lane; k < G; k += 32) s += a[k];
        const float mean = warp_sum(s) / (float)G;
        float q = 0.f;
        for (int k = lane; k < G; k += 32) { const float d = a[k] - mean; q += d * d; }
        const float rstd = rsqrtf(warp_sum(q) / (float)G + kEps);
        const size_t off = (size_t)(row0 + r) * G;
        for (int k = lane; k < G; k += 32) {
            const float xh = (a[k] - mean) * rstd;
            const float y = xh * gamma[k] + beta[k];
            a[k] = y;
            if (xhat_out) { xhat_out[off + k] = xh; y_out[off + k] = y; }
        }
        if (rstd_out && lane == 0) rstd_out[row0 + r] = rstd;
    }
}

__global__ void __launch_bounds__(kThreads) global_post_fwd_kernel(PostFwd p) {
    extern __shared__ __align__(16) float sm[];
    __shared__ float red[32];
    __shared__ int flag_s;
    const int G = p.G, MT = row_tile_of(G), lda = G + 1;
    float* A_s = sm;                                   // [MT][lda]  g1 rows
    float* Bt_s = A_s + (size_t)MT * lda;              // [G][8]
    float* red_s = Bt_s + (size_t)G * kNT;             // [256*8]
    const int n0 = blockIdx.x * kNT;
    const float coef = coef_of(p.Wp, p.Kdim, red);
    load_bt_rows(Bt_s, p.W4, G, n0, min(kNT, G - n0), G);
    const bool writer = blockIdx.x == 0;
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();
        // LayerNorm 1 of every row of the tile, recomputed by every CTA (CTA 0 also publishes g1 / xhat1 / rstd1)
        fill_tile<3>(A_s, lda, MT, mt, G,
                     [&](int r, int k, int i) { return ld4((i == 0 ? p.g : (i == 1 ? p.gh : p.pool)) + (size_t)(row0 + r) * G + k); },
                     [&](float4 (&v)[3], int, int) {
                         return make_float4(v[0].x + v[1].x + coef * v[2].x, v[0].y + v[1].y + coef * v[2].y,
                                            v[0].z + v[1].z + coef * v[2].z, v[0].w + v[1].w + coef * v[2].w);
                     });
        __syncthreads();
        ln_rows_smem(A_s, lda, mt, G, p.gamma1, p.beta1, writer ? p.xhat1 : nullptr, p.g1, writer ? p.rstd1 : nullptr, row0);
        __syncthreads();
        float out[kNT];
        cta_gemm8(A_s, lda, Bt_s, G, MT, red_s, out);
        if (threadIdx.x < mt) {
            float* o = p.z4 + (size_t)(row0 + threadIdx.x) * G + n0;
#pragma unroll
            for (int j = 0; j < kNT; ++j)
                if (n0 + j < G) o[j] = out[j] + p.b4[n0 + j];
        }
    }
    // the last CTA to finish normalises g1 + gelu(z4) for all rows
    if (!last_cta_done(p.counter, &flag_s)) return;
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();
        fill_tile<2>(A_s, lda, MT, mt, G,
                     [&](int r, int k, int i) { return ld4cg((i == 0 ? p.g1 : p.z4) + (size_t)(row0 + r) * G + k); },
                     [&](float4 (&v)[2], int, int) {
                         return make_float4(v[0].x + gelu_f(v[1].x), v[0].y + gelu_f(v[1].y), v[0].z + gelu_f(v[1].z),
                                            v[0].w + gelu_f(v[1].w));
                     });
        __syncthreads();
        ln_rows_smem(A_s, lda, mt, G, p.gamma2, p.beta2, p.xhat2, p.g2, p.rstd2, row0);
    }
}

// ============================================================================================================
// post backward (row tiles of 32: two shared arrays, the gradient rows and the matching xhat rows)
// ============================================================================================================
struct PostBwd {
    const float *dg2, *gamma2, *xhat2, *rstd2, *z4, *g1, *W4, *gamma1, *xhat1, *rstd1, *pool, *Wp;
    float *dg1, *dx1, *dpool, *dWp, *dgamma1, *dbeta1, *dW4, *db4, *dgamma2, *dbeta2;
    int* counter;
    int B, G, Kdim;
};
constexpr int kMT2 = 32;

// per-row means m1 = mean(d), m2 = mean(d * xh) over a shared tile (d already includes gamma when with_gamma == nullptr)
__device__ __forceinline__ void row_means(const float* A_s, const float* X_s, int lda, int mt, int G, const float* __restrict__ gamma,
                                          float* m_s) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < mt; r += kThreads / 32) {
        const float* a = A_s + (size_t)r * lda;
        const float* x = X_s + (size_t)r * lda;
        float s1 = 0.f, s2 = 0.f;
        for (int k = lane; k < G; k += 32) {
            const float d = gamma ? a[k] * gamma[k] : a[k];
            s1 += d;
            s2 += d * x[k];
        }
        s1 = warp_sum(s1); s2 = warp_sum(s2);
        if (lane == 0) { m_s[2 * r] = s1 / (float)G; m_s[2 * r + 1] = s2 / (float)G; }
    }
}

__global__ void __launch_bounds__(kThreads) global_post_bwd_kernel(PostBwd p) {
    extern __shared__ __align__(16) float sm[];
    __shared__ float red[32];
    __shared__ int flag_s;
    const int G = p.G, MT = kMT2, lda = G + 1;
    float* A_s = sm;                                   // [32][lda]  d = dg2*gamma2, later dz4
    float* X2_s = A_s + (size_t)MT * lda;              // [32][lda]  xhat rows
    float* Bt_s = X2_s + (size_t)MT * lda;             // [G][8]     W4[:, n0..n0+7]
    float* red_s = Bt_s + (size_t)G * kNT;             // [256*8]
    float* D_s = red_s + kThreads * kNT;               // [32][8] dz4 slice
    float* Xs_s = D_s + MT * kNT;                      // [32][8] dx2 slice (residual path)
    float* m_s = Xs_s + MT * kNT;                      // [32][2]
    const int n0 = blockIdx.x * kNT;
    load_bt_cols(Bt_s, p.W4, G, n0, min(kNT, G - n0), G);
    float accw[kQ][kNT];
#pragma unroll
    for (int q = 0; q < kQ; ++q)
#pragma unroll
        for (int j = 0; j < kNT; ++j) accw[q][j] = 0.f;
    float c_db4 = 0.f, c_dg2 = 0.f, c_db2 = 0.f;       // column sums for column n0 + threadIdx.x (< 8)
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();
        fill_tile<1>(A_s, lda, MT, mt, G,
                     [&](int r, int k, int) { return ld4(p.dg2 + (size_t)(row0 + r) * G + k); },
                     [&](float4 (&v)[1], int, int k) {
                         const float4 gm = ld4(p.gamma2 + k);
                         return make_float4(v[0].x * gm.x, v[0].y * gm.y, v[0].z * gm.z, v[0].w * gm.w);
                     });
        fill_tile<1>(X2_s, lda, MT, mt, G,
                     [&](int r, int k, int) { return ld4(p.xhat2 + (size_t)(row0 + r) * G + k); },
                     [&](float4 (&v)[1], int, int) { return v[0]; });
        __syncthreads();
        row_means(A_s, X2_s, lda, mt, G, nullptr, m_s);
        __syncthreads();
        // dx2 = rstd2*(d - m1 - xhat2*m2); keep own columns of dx2; A <- dz4 = dx2 * gelu'(z4)
        fill_tile<1>(A_s, lda, MT, mt, G,
                     [&](int r, int k, int) { return ld4(p.z4 + (size_t)(row0 + r) * G + k); },
                     [&](float4 (&v)[1], int r, int k) {
                         const float rs = p.rstd2[row0 + r], m1 = m_s[2 * r], m2 = m_s[2 * r + 1];
                         const float* a = A_s + (size_t)r * lda + k;
                         const float* x = X2_s + (size_t)r * lda + k;
                         const float d0 = rs * (a[0] - m1 - x[0] * m2), d1 = rs * (a[1] - m1 - x[1] * m2);
                         const float d2 = rs * (a[2] - m1 - x[2] * m2), d3 = rs * (a[3] - m1 - x[3] * m2);
                         if (k >= n0 && k < n0 + kNT) {
                             float* xs = Xs_s + r * kNT + (k - n0);
                             xs[0] = d0; xs[1] = d1; xs[2] = d2; xs[3] = d3;
                         }
                         return make_float4(d0 * gelu_grad_f(v[0].x), d1 * gelu_grad_f(v[0].y), d2 * gelu_grad_f(v[0].z),
                                            d3 * gelu_grad_f(v[0].w));
                     });
        __syncthreads();
        for (int i = threadIdx.x; i < MT * kNT; i += kThreads) {
            const int m = i / kNT, j = i - m * kNT;
            D_s[i] = (m < mt && n0 + j < G) ? A_s[(size_t)m * lda + n0 + j] : 0.f;
        }
        __syncthreads();
        if (threadIdx.x < kNT && n0 + threadIdx.x < G) {
            const int c = n0 + threadIdx.x;
            for (int m = 0; m < mt; ++m) {
                const size_t o = (size_t)(row0 + m) * G + c;
                c_db4 += D_s[m * kNT + threadIdx.x];
                c_dg2 += p.dg2[o] * p.xhat2[o];
                c_db2 += p.dg2[o];
            }
        }
        dw_accumulate<kQ>(accw, D_s, p.g1, G, row0, mt, G);             // dW4[n0.., :] += dz4[:, n0..]^T g1
        float out[kNT];
        cta_gemm8(A_s, lda, Bt_s, G, MT, red_s, out);                   // dz4 W4[:, n0..]
        if (threadIdx.x < mt) {
            float* o = p.dg1 + (size_t)(row0 + threadIdx.x) * G + n0;
#pragma unroll
            for (int j = 0; j < kNT; ++j)
                if (n0 + j < G) o[j] = out[j] + Xs_s[threadIdx.x * kNT + j];
        }
    }
    dw_store<kQ>(accw, p.dW4, G, n0, min(kNT, G - n0), G);
    if (threadIdx.x < kNT && n0 + threadIdx.x < G) {
        p.db4[n0 + threadIdx.x] = c_db4;
        p.dgamma2[n0 + threadIdx.x] = c_dg2;
        p.dbeta2[n0 + threadIdx.x] = c_db2;
    }
    // the last CTA runs LayerNorm-1 backward over complete rows of dg1
    if (!last_cta_done(p.counter, &flag_s)) return;
    const float coef = coef_of(p.Wp, p.Kdim, red);
    float cg[kQ], cb[kQ];                              // dgamma1 / dbeta1 of columns threadIdx.x + q*256
#pragma unroll
    for (int q = 0; q < kQ; ++q) { cg[q] = 0.f; cb[q] = 0.f; }
    float dcoef = 0.f;
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();
        fill_tile<1>(A_s, lda, MT, mt, G, [&](int r, int k, int) { return ld4cg(p.dg1 + (size_t)(row0 + r) * G + k); },
                     [&](float4 (&v)[1], int, int) { return v[0]; });
        fill_tile<1>(X2_s, lda, MT, mt, G, [&](int r, int k, int) { return ld4(p.xhat1 + (size_t)(row0 + r) * G + k); },
                     [&](float4 (&v)[1], int, int) { return v[0]; });
        __syncthreads();
        row_means(A_s, X2_s, lda, mt, G, p.gamma1, m_s);
        __syncthreads();
        // column-wise: thread walks the tile's rows for its columns (shared memory, conflict-free)
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            const int k = threadIdx.x + q * kThreads;
            if (k < G)
                for (int r = 0; r < mt; ++r) {
                    const float d = A_s[(size_t)r * lda + k];
                    cg[q] += d * X2_s[(size_t)r * lda + k];
                    cb[q] += d;
                }
        }
        // element-wise: dx1, dpool, and the W_parameter coefficient
        const int kv = G / 4, total = mt * kv;
        for (int i = threadIdx.x; i < total; i += kThreads) {
            const int r = i / kv, k = (i - r * kv) * 4;
            const size_t o = (size_t)(row0 + r) * G + k;
            const float4 pl = ld4(p.pool + o), gm = ld4(p.gamma1 + k);
            const float rs = p.rstd1[row0 + r], m1 = m_s[2 * r], m2 = m_s[2 * r + 1];
            const float* a = A_s + (size_t)r * lda + k;
            const float* x = X2_s + (size_t)r * lda + k;
            float4 dx;
            dx.x = rs * (a[0] * gm.x - m1 - x[0] * m2); dx.y = rs * (a[1] * gm.y - m1 - x[1] * m2);
            dx.z = rs * (a[2] * gm.z - m1 - x[2] * m2); dx.w = rs * (a[3] * gm.w - m1 - x[3] * m2);
            *reinterpret_cast<float4*>(p.dx1 + o) = dx;
            *reinterpret_cast<float4*>(p.dpool + o) = make_float4(coef * dx.x, coef * dx.y, coef * dx.z, coef * dx.w);
            dcoef += dx.x * pl.x + dx.y * pl.y + dx.z * pl.z + dx.w * pl.w;
        }
    }
#pragma unroll
    for (int q = 0; q < kQ; ++q) {
        const int k = threadIdx.x + q * kThreads;
        if (k < G) { p.dgamma1[k] = cg[q]; p.dbeta1[k] = cb[q]; }
    }
    const float tot = block_sum(dcoef, red) / (float)p.Kdim;
    for (int k = threadIdx.x; k < p.Kdim; k += kThreads) p.dWp[k] = tot;
}

size_t smem_bytes(int MT, int KA, bool with_slices) {
    size_t f = (size_t)MT * (KA + 1) + (size_t)KA * kNT + (size_t)kThreads * kNT;
    if (with_slices) f += 2 * (size_t)MT * kNT;
    return f * sizeof(float) + 64;
}
size_t smem_bytes_post_bwd(int G) {
    const size_t f = 2 * (size_t)kMT2 * (G + 1) + (size_t)G * kNT + (size_t)kThreads * kNT + 2 * (size_t)kMT2 * kNT + 2 * kMT2;
    return f * sizeof(float) + 64;
}

template <class Kern>
int set_smem(Kern kern, size_t bytes) {
    if (bytes > 48 * 1024) PB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return PB_OK;
}

}  // namespace

extern "C" {

int pb_global_fused_ok(int B, int G, int C) {
    if (B < 1 || G < 8 || C < 8 || G % 8 || C % 8) return 0;
    if (G > 1024 || G + C > 2048) return 0;                       // dW rows: up to kQ*256 columns per thread sweep
    if (G % 4 || C % 4) return 0;
    if (smem_bytes(host_row_tile(G + C), G + C, true) > 220 * 1024) return 0;
    if (smem_bytes_post_bwd(G) > 220 * 1024) return 0;
    return 1;
}

int pb_global_pre_fwd(const float* g, const float* W1, const float* b1, const float* W3, const float* b3, float* zg, float* g2l,
                      float* z3, float* gh, int B, int G, int C, void* stream) {
    PB_REQUIRE(pb_global_fused_ok(B, G, C), PB_ERR_UNSUPPORTED, "global_pre_fwd: shape B=%d G=%d C=%d not covered", B, G, C);
    const size_t smem = smem_bytes(host_row_tile(G), G, false);
    int rc = set_smem(global_pre_fwd_kernel, smem);
    if (rc) return rc;
    PreFwd p{g, W1, b1, W3, b3, zg, g2l, z3, gh, B, G, C};
    global_pre_fwd_kernel<<<(G + C) / kNT, kThreads, smem, (cudaStream_t)stream>>>(p);
    pb_count_launch(1);
    return pb_check_launch("pb_global_pre_fwd");
}

int pb_global_pre_bwd(const float* dg2l, const float* dgh, const float* zg, const float* z3, const float* g, const float* W1,
                      const float* W3, float* dg, float* dW1, float* db1, float* dW3, float* db3, int B, int G, int C,
                      void* stream) {
    PB_REQUIRE(pb_global_fused_ok(B, G, C), PB_ERR_UNSUPPORTED, "global_pre_bwd: shape B=%d G=%d C=%d not covered", B, G, C);
    const size_t smem = smem_bytes(host_row_tile(G + C), G + C, true);
    int rc = set_smem(global_pre_bwd_kernel, smem);
    if (rc) return rc;
    PreBwd p{dg2l, dgh, zg, z3, g, W1, W3, dg, dW1, db1, dW3, db3, B, G, C};
    global_pre_bwd_kernel<<<(G + C) / kNT, kThreads, smem, (cudaStream_t)stream>>>(p);
    pb_count_launch(1);
    return pb_check_launch("pb_global_pre_bwd");
}

int pb_global_post_fwd(const float* g, const float* gh, const float* pool, const float* Wp, int Kdim, const float* gamma1,
                       const float* beta1, const float* W4, const float* b4, const float* gamma2, const float* beta2, float* g1,
                       float* xhat1, float* rstd1, float* z4, float* xhat2, float* rstd2, float* g2, int* counter, int B, int G,
                       void* stream) {
    PB_REQUIRE(pb_global_fused_ok(B, G, 8), PB_ERR_UNSUPPORTED, "global_post_fwd: shape B=%d G=%d not covered", B, G);
    PB_REQUIRE(counter != nullptr && Kdim > 0, PB_ERR_SHAPE, "global_post_fwd: counter / Kdim missing");
    const size_t smem = smem_bytes(host_row_tile(G), G, false);
    int rc = set_smem(global_post_fwd_kernel, smem);
    if (rc) return rc;
    PostFwd p{g, gh, pool, Wp, gamma1, beta1, W4, b4, gamma2, beta2, g1, xhat1, rstd1, z4, xhat2, rstd2, g2, counter, B, G, Kdim};
    global_post_fwd_kernel<<<G / kNT, kThreads, smem, (cudaStream_t)stream>>>(p);
    pb_count_launch(1);
    return pb_check_launch("pb_global_post_fwd");
}

int pb_global_post_bwd(const float* dg2, const float* gamma2, const float* xhat2, const float* rstd2, const float* z4,
                       const float* g1, const float* W4, const float* gamma1, const float* xhat1, const float* rstd1,
                       const float* pool, const float* Wp, int Kdim, float* dg1_buf, float* dx1, float* dpool, float* dWp,
                       float* dgamma1, float* dbeta1, float* dW4, float* db4, float* dgamma2, float* dbeta2, int* counter, int B,
                       int G, void* stream) {
    PB_REQUIRE(pb_global_fused_ok(B, G, 8), PB_ERR_UNSUPPORTED, "global_post_bwd: shape B=%d G=%d not covered", B, G);
    PB_REQUIRE(counter != nullptr && Kdim > 0, PB_ERR_SHAPE, "global_post_bwd: bad arguments");
    const size_t smem = smem_bytes_post_bwd(G);
    int rc = set_smem(global_post_bwd_kernel, smem);
    if (rc) return rc;
    PostBwd p{dg2, gamma2, xhat2, rstd2, z4, g1, W4, gamma1, xhat1, rstd1, pool, Wp,
              dg1_buf, dx1, dpool, dWp, dgamma1, dbeta1, dW4, db4, dgamma2, dbeta2, counter, B, G, Kdim};
    global_post_bwd_kernel<<<G / kNT, kThreads, smem, (cudaStream_t)stream>>>(p);
    pb_count_launch(1);
    return pb_check_launch("pb_global_post_bwd");
}

}  // extern "C"
