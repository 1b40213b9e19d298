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
// Every kernel alternates between two decompositions of the same [B x G] problem:
//   row phases    (LayerNorm forward/backward): one batch row per CTA, all 512 threads across its columns;
//   column phases (the GEMMs): each CTA owns 8 output columns, stages all rows of the A operand in shared memory
//                 ([rows x K] fp32) and contracts it against its [K x 8] weight slice.
// Phases are separated by a cooperative grid barrier (the grids are <= 148 CTAs of one wave, launched with
// cudaLaunchCooperativeKernel), so the row-wise results never have to be recomputed per column slice.
//
// Two things bound these kernels in round 1 (profiles/r01_global_phases.md): staging the shared A operand with
// LDG + scalar STS (3.8 us for 128 KB per CTA) and the scalar-FMA 64 x 8 x K product (1.7 - 5 us).  Now
//   * operands that are plain copies arrive by TMA bulk copies (cp.async.bulk, one per row into a row pitch of K + 4
//     floats, completion on an mbarrier): no register round trip, no store-side bank conflicts, and the copies of
//     weights that do not depend on the previous kernel are issued before griddepcontrol.wait;
//   * the product runs on the tensor cores through warp-level mma.sync.m16n8k8 TF32 with the error-compensated
//     three-product split (a = hi + lo, b = hi + lo; hi*hi + hi*lo + lo*hi), which keeps fp32-level accuracy
//     (relative error ~2^-22 per product): 16 warps = (row tiles of 16) x (K groups), partials reduced through shared
//     memory.  The pitch K + 4 makes the A-fragment reads (8 rows x 4 consecutive floats) bank-conflict free.
#include <cooperative_groups.h>
#include "tc_common.cuh"
#include "../../include/pb_b200.h"

namespace cg = cooperative_groups;

void* pb_debug_ptr();          // tc_host.cu: optional device buffer for phase timestamps (scripts/global_phase_probe.py)

namespace {

constexpr int kNT = 8;            // output columns per CTA
constexpr int kThreads = 512;
constexpr float kEps = 1e-5f;
constexpr int kQ = 2;             // row phases / dW row sweep: up to kQ * 512 = 1024 columns

constexpr int kPad = 4;           // A tile row pitch = K + kPad floats: 16-byte aligned rows, conflict-free mma fragment reads

#ifdef PB_GLOBAL_PROBE            // phase timestamps for scripts/global_phase_probe.py: compile-time only
__device__ __forceinline__ void stamp(unsigned long long* dbg, int slot) {
    if (dbg != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        dbg[slot] = t;
    }
}
#else
__device__ __forceinline__ void stamp(unsigned long long*, int) {}
#endif

// ---- TMA bulk row copies: rows [row0, row0 + mt) of a row-major [*, ld] fp32 tensor -> A_s rows of pitch lda ----------
// Called by ONE warp.  K * 4 bytes per row (K % 4 == 0), source rows 16-byte aligned (ld % 4 == 0).  Completion: `bar`
// receives mt * K * 4 bytes; the caller waits on it with the barrier's current parity.
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_rows(float* A_s, int lda, const float* src, size_t ld, int row0, int mt, int K, uint64_t* bar) {
    const int lane = threadIdx.x & 31;
    if (lane == 0) mbar_expect_tx(bar, (uint32_t)mt * (uint32_t)K * 4u);
    __syncwarp();
    for (int r = lane; r < mt; r += 32) bulk_copy_g2s(A_s + (size_t)r * lda, src + (size_t)(row0 + r) * ld, (uint32_t)K * 4u, bar);
}
__device__ __forceinline__ void zero_rows(float* A_s, int lda, int mt, int MT, int K) {      // rows mt .. MT-1 (ragged last tile)
    const int kv = K / 4;
    for (int i = threadIdx.x; i < (MT - mt) * kv; i += blockDim.x) {
        const int r = mt + i / kv, k = (i % kv) * 4;
        *reinterpret_cast<float4*>(A_s + (size_t)r * lda + k) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

__host__ __device__ __forceinline__ int row_tile_of(int K) { return K <= 704 ? 64 : (K <= 1472 ? 32 : 16); }

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 ld4cg(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// ---- CTA-level product: out[r][j] = sum_k A_s[r*lda + k] * Bt_s[k*8 + j],  r < MT, j < 8, k < Kd -----------------
// Tensor cores, warp-level: mma.sync.m16n8k8 TF32 x 3 (error-compensated split, see the file header).  Warp w owns row
// tile w % (MT/16) and K group w / (MT/16); the K groups' partial tiles are combined through red_s [groups][MT][8].
// The result is returned in out[] for threads tid < MT (row tid).  MT in {16, 32, 64}, Kd % 8 == 0, lda % 32 == 4.
__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void cta_gemm8(const float* A_s, int lda, const float* Bt_s, int Kd, int MT, float* red_s,
                                          float (&out)[kNT]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int nmt = MT >> 4, nkg = (kThreads / 32) / nmt;
    const int mtile = warp % nmt, kg = warp / nmt;
    const int steps = Kd >> 3, s0 = kg * steps / nkg, s1 = (kg + 1) * steps / nkg;
    float c[4] = {0.f, 0.f, 0.f, 0.f};
    const float* a0p = A_s + (size_t)(mtile * 16 + g) * lda + t;
    const float* a1p = a0p + 8 * lda;
#pragma unroll 2
    for (int s = s0; s < s1; ++s) {
        const int k = s << 3;
        const float af[4] = {a0p[k], a1p[k], a0p[k + 4], a1p[k + 4]};
        const float bf[2] = {Bt_s[(k + t) * kNT + g], Bt_s[(k + t + 4) * kNT + g]};
        uint32_t ah[4], al[4], bh[2], bl[2];
#pragma unroll
        for (int i = 0; i < 4; ++i) { ah[i] = to_tf32(af[i]); al[i] = to_tf32(af[i] - __uint_as_float(ah[i])); }
#pragma unroll
        for (int i = 0; i < 2; ++i) { bh[i] = to_tf32(bf[i]); bl[i] = to_tf32(bf[i] - __uint_as_float(bh[i])); }
        mma_tf32(c, al, bh);                   // small terms first
        mma_tf32(c, ah, bl);
        mma_tf32(c, ah, bh);
    }
    {
        float* r0 = red_s + ((size_t)kg * MT + mtile * 16 + g) * kNT + 2 * t;
        r0[0] = c[0]; r0[1] = c[1];
        r0[8 * kNT] = c[2]; r0[8 * kNT + 1] = c[3];
    }
    __syncthreads();
    if (threadIdx.x < MT) {
#pragma unroll
        for (int j = 0; j < kNT; ++j) out[j] = 0.f;
        for (int q = 0; q < nkg; ++q) {
            const float4 v0 = *reinterpret_cast<const float4*>(red_s + ((size_t)q * MT + threadIdx.x) * kNT);
            const float4 v1 = *reinterpret_cast<const float4*>(red_s + ((size_t)q * MT + threadIdx.x) * kNT + 4);
            out[0] += v0.x; out[1] += v0.y; out[2] += v0.z; out[3] += v0.w;
            out[4] += v1.x; out[5] += v1.y; out[6] += v1.z; out[7] += v1.w;
        }
    }
    __syncthreads();
}

// Bt_s[k][j] = W[(n0 + j) * ldw + k]   (weight rows are the output columns: "X W^T").  Kd % 4 == 0, ldw % 4 == 0.
__device__ __forceinline__ void load_bt_rows(float* Bt_s, const float* __restrict__ W, int ldw, int n0, int Kd) {
    const int kv = Kd / 4, total = kNT * kv;
#pragma unroll 2
    for (int i = threadIdx.x; i < total; i += kThreads) {
        const int j = i / kv, k = (i - j * kv) * 4;
        const float4 v = ld4(W + (size_t)(n0 + j) * ldw + k);
        Bt_s[(k + 0) * kNT + j] = v.x; Bt_s[(k + 1) * kNT + j] = v.y;
        Bt_s[(k + 2) * kNT + j] = v.z; Bt_s[(k + 3) * kNT + j] = v.w;
    }
}
// Bt_s[k][j] = W[k * ldw + c0 + j]     (weight columns are the output columns: "dZ W").  c0 % 8 == 0, ldw % 4 == 0.
__device__ __forceinline__ void load_bt_cols(float* Bt_s, const float* __restrict__ W, int ldw, int c0, int Kd) {
    const int total = Kd * 2;
#pragma unroll 2
    for (int i = threadIdx.x; i < total; i += kThreads) {
        const int k = i >> 1, h = (i & 1) * 4;
        *reinterpret_cast<float4*>(Bt_s + k * kNT + h) = ld4(W + (size_t)k * ldw + c0 + h);
    }
}

// Cooperative copy of rows [row0, row0 + mt) of a row-major [*, K] tensor into the shared tile (pitch lda), with an
// optional element-wise transform of up to two sources; rows >= mt are zero-filled.  kCG: bypass L1 (data written
// earlier in this kernel by other CTAs).
template <int NSRC, bool kCG, class Combine>
__device__ __forceinline__ void fill_tile(float* A_s, int lda, int MT, int mt, int K, const float* s0, const float* s1, size_t ld,
                                          int row0, Combine combine) {
    const int kv = K / 4, total = MT * kv;
    constexpr int U = NSRC == 1 ? 8 : 4;
    for (int base = threadIdx.x; base < total; base += kThreads * U) {
        float4 v[U][2];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * kThreads;
            const int r = i / kv, k = (i - r * kv) * 4;
            if (i < total && r < mt) {
                const size_t o = (size_t)(row0 + r) * ld + k;
                v[u][0] = kCG ? ld4cg(s0 + o) : ld4(s0 + o);
                if (NSRC == 2) v[u][1] = kCG ? ld4cg(s1 + o) : ld4(s1 + o);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * kThreads;
            if (i < total) {
                const int r = i / kv, k = (i - r * kv) * 4;
                float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < mt) o = combine(v[u][0], v[u][1]);
                *reinterpret_cast<float4*>(A_s + (size_t)r * lda + k) = o;
            }
        }
    }
}

// Weight gradients from the staged tile: the A tile in shared memory already holds every row of dZ for ALL its columns,
// so a CTA can produce dW[n, k0 .. k0+7] for every n from it and an 8-column slice of X (64 x 8 floats) -- instead of
// sweeping all rows of X from L2 for its own 8 rows of dW.  acc[q][j] += sum_m A_s[m][n] * Xs[m][j], n = tid + q*512.
__device__ __forceinline__ void fill_x_slice(float* Xs, const float* __restrict__ X, int ldx, int row0, int mt, int MT, int k0) {
    for (int i = threadIdx.x; i < MT * kNT; i += kThreads) {
        const int m = i / kNT, j = i - m * kNT;
        Xs[i] = m < mt ? __ldg(X + (size_t)(row0 + m) * ldx + k0 + j) : 0.f;
    }
}
__device__ __forceinline__ void dw_from_tile(float (&acc)[kQ][kNT], const float* A_s, int lda, const float* Xs, int mt, int Nrows) {
#pragma unroll 4
    for (int m = 0; m < mt; ++m) {
        const float4 x0 = *reinterpret_cast<const float4*>(Xs + m * kNT);
        const float4 x1 = *reinterpret_cast<const float4*>(Xs + m * kNT + 4);
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            const int n = threadIdx.x + q * kThreads;
            if (n < Nrows) {
                const float a = A_s[(size_t)m * lda + n];
                acc[q][0] = fmaf(a, x0.x, acc[q][0]); acc[q][1] = fmaf(a, x0.y, acc[q][1]);
                acc[q][2] = fmaf(a, x0.z, acc[q][2]); acc[q][3] = fmaf(a, x0.w, acc[q][3]);
                acc[q][4] = fmaf(a, x1.x, acc[q][4]); acc[q][5] = fmaf(a, x1.y, acc[q][5]);
                acc[q][6] = fmaf(a, x1.z, acc[q][6]); acc[q][7] = fmaf(a, x1.w, acc[q][7]);
            }
        }
    }
}
__device__ __forceinline__ void store8(float* dst, const float (&v)[kNT]) {
    *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ float coef_of(const float* __restrict__ Wp, int Kdim, float* red /* >= 32 */) {
    float t = 0.f;
    for (int k = threadIdx.x; k < Kdim; k += kThreads) t += Wp[k];
    return block_sum(t, red) / (float)Kdim;
}

struct MulGeluGrad4 {
    __device__ float4 operator()(const float4 d, const float4 z) const {
        return make_float4(d.x * gelu_grad_f(z.x), d.y * gelu_grad_f(z.y), d.z * gelu_grad_f(z.z), d.w * gelu_grad_f(z.w));
    }
};
struct Copy4 {
    __device__ float4 operator()(const float4 a, const float4) const { return a; }
};

// ============================================================================================================
// pre forward:  [zg | z3] = g [W1 ; W3]^T + [b1 ; b3];  g2l = gelu(zg), gh = gelu(z3)      (no row phase)
// ============================================================================================================
struct PreFwd {
    const float *g, *W1, *b1, *W3, *b3;
    float *zg, *g2l, *z3, *gh;
    int B, G, C;
};

__global__ void __launch_bounds__(kThreads, 1) global_pre_fwd_kernel(PreFwd p) {
    extern __shared__ __align__(16) float sm[];
    __shared__ __align__(8) uint64_t bar;
    const int K = p.G, MT = row_tile_of(K), lda = K + kPad;
    float* A_s = sm;                                   // [MT][lda]
    float* Bt_s = A_s + (size_t)MT * lda;              // [K][8]
    float* red_s = Bt_s + (size_t)K * kNT;             // [16 warps' partial tiles] <= [512*8]
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    const int n0 = blockIdx.x * kNT;                   // column in the concatenated [C + G] output
    const bool is_g2l = n0 < p.C;                      // slices never straddle (C % 8 == 0)
    const float* W = is_g2l ? p.W1 : p.W3;
    const float* bias = is_g2l ? p.b1 : p.b3;
    const int c0 = is_g2l ? n0 : n0 - p.C;
    const int ncols = is_g2l ? p.C : p.G;
    load_bt_rows(Bt_s, W, K, c0, K);                   // weights do not depend on the previous kernel: before the wait
    __syncthreads();
    pdl_wait();
    uint32_t phase = 0;
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();                               // the previous tile's product has finished reading A_s
        if (threadIdx.x < 32) tma_rows(A_s, lda, p.g, K, row0, mt, K, &bar);
        zero_rows(A_s, lda, mt, MT, K);
        mbar_wait(&bar, phase);
        phase ^= 1;
        __syncthreads();
        float out[kNT];
        cta_gemm8(A_s, lda, Bt_s, K, MT, red_s, out);
        if (threadIdx.x < mt) {
            const int row = row0 + threadIdx.x;
            float* zo = (is_g2l ? p.zg : p.z3) + (size_t)row * ncols + c0;
            float* yo = (is_g2l ? p.g2l : p.gh) + (size_t)row * ncols + c0;
            float z[kNT], y[kNT];
#pragma unroll
            for (int j = 0; j < kNT; ++j) { z[j] = out[j] + bias[c0 + j]; y[j] = gelu_f(z[j]); }
            *reinterpret_cast<float4*>(zo) = make_float4(z[0], z[1], z[2], z[3]);
            *reinterpret_cast<float4*>(zo + 4) = make_float4(z[4], z[5], z[6], z[7]);
            *reinterpret_cast<float4*>(yo) = make_float4(y[0], y[1], y[2], y[3]);
            *reinterpret_cast<float4*>(yo + 4) = make_float4(y[4], y[5], y[6], y[7]);
        }
    }
}

// ============================================================================================================
// pre backward: dzg = dg2l*gelu'(zg), dz3 = dgh*gelu'(z3);  dg = dz3 W3 + dzg W1;  dW3 = dz3^T g, dW1 = dzg^T g
// (the activation derivative is element-wise, so it is applied while the A tile is staged: no row phase)
// ============================================================================================================
struct PreBwd {
    const float *dg2l, *dgh, *zg, *z3, *g, *W1, *W3;
    float *dg, *dW1, *db1, *dW3, *db3;
    int B, G, C;
    int dgh_is_dz3;                // dgh already holds dgh * gelu'(z3) (written by the post backward kernel)
    float* dzg_out;                // optional [B, C]: dg2l * gelu'(zg), for a caller that forms dW1 itself
    int skip_dw;                   // leave dW3 / dW1 to the caller (pb_matmul_tn on another stream)
    unsigned long long* dbg;
};

template <bool kDz3Ready>
__global__ void __launch_bounds__(kThreads, 1) global_pre_bwd_kernel(PreBwd p) {
    extern __shared__ __align__(16) float sm[];
    const int G = p.G, C = p.C, KA = G + C, MT = row_tile_of(KA), lda = KA + kPad;
    float* A_s = sm;                                   // [MT][lda] = [dz3 | dzg]
    float* Bt_s = A_s + (size_t)MT * lda;              // [KA][8]
    float* red_s = Bt_s + (size_t)KA * kNT;            // [512*8]
    float* D_s = red_s + kThreads * kNT;               // [MT][8] slice g[:, n0 .. n0+7] for the dW product
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    uint32_t phase = 0;
    // CTA i owns columns [8i, 8i+8) of dg and of the stacked [dW3 ; dW1] matrix (if 8i < G), and db entries [8i, 8i+8).
    const int n0 = blockIdx.x * kNT;                   // 0 .. G + C
    const bool has_dg = n0 < G;
    stamp(p.dbg, 16);
    if (has_dg) {                                      // Bt = [W3[:, n0..] ; W1[:, n0..]]  (rows = contraction index)
        load_bt_cols(Bt_s, p.W3, G, n0, G);            // weights: before the wait on the previous kernel
        load_bt_cols(Bt_s + (size_t)G * kNT, p.W1, G, n0, C);
    }
    __syncthreads();
    pdl_wait();
    stamp(p.dbg, 22);
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
        if (kDz3Ready) {                               // dz3 is a plain copy: TMA rows, zero the ragged tail
            if (threadIdx.x < 32) tma_rows(A_s, lda, p.dgh, G, row0, mt, G, &bar);
            zero_rows(A_s, lda, mt, MT, G);
        } else {
            fill_tile<2, false>(A_s, lda, MT, mt, G, p.dgh, p.z3, G, row0, MulGeluGrad4());
        }
        stamp(p.dbg, 23);
        fill_tile<2, false>(A_s + G, lda, MT, mt, C, p.dg2l, p.zg, C, row0, MulGeluGrad4());
        if (kDz3Ready) { mbar_wait(&bar, phase); phase ^= 1; }
        __syncthreads();
        stamp(p.dbg, 17);
        if (p.dzg_out != nullptr && blockIdx.x == 0)       // one CTA publishes dzg (columns G .. G+C of the staged tile)
            for (int i = threadIdx.x; i < mt * C; i += kThreads) p.dzg_out[(size_t)(row0 + i / C) * C + i % C] = A_s[(size_t)(i / C) * lda + G + i % C];
        if (has_dg && !p.skip_dw) fill_x_slice(D_s, p.g, G, row0, mt, MT, n0);         // g[:, n0 .. n0+7]
        if (threadIdx.x < 32) {                        // db of columns n0 .. n0+7 of [dz3 | dzg] (n0 + j < KA always)
            float a = 0.f;
            for (int m = threadIdx.x >> 3; m < mt; m += 4) a += A_s[(size_t)m * lda + n0 + (threadIdx.x & 7)];
            a += __shfl_xor_sync(0xffffffffu, a, 8); a += __shfl_xor_sync(0xffffffffu, a, 16);
            colsum += a;                               // read by threads < 8 only
        }
        __syncthreads();
        stamp(p.dbg, 18);
        if (has_dg && !p.skip_dw) dw_from_tile(accw, A_s, lda, D_s, mt, KA);          // [dW3 ; dW1][:, n0 .. n0+7]
        stamp(p.dbg, 19);
        if (has_dg) {
            float out[kNT];
            cta_gemm8(A_s, lda, Bt_s, KA, MT, red_s, out);
            if (threadIdx.x < mt) {
                float* o = p.dg + (size_t)(row0 + threadIdx.x) * G + n0;
                *reinterpret_cast<float4*>(o) = make_float4(out[0], out[1], out[2], out[3]);
                *reinterpret_cast<float4*>(o + 4) = make_float4(out[4], out[5], out[6], out[7]);
            }
        }
    }
    stamp(p.dbg, 20);
    if (has_dg && !p.skip_dw) {
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            const int n = threadIdx.x + q * kThreads;
            if (n < KA) store8((n < G ? p.dW3 + (size_t)n * G : p.dW1 + (size_t)(n - G) * G) + n0, accw[q]);
        }
    }
    if (threadIdx.x < kNT) (w_is_3 ? p.db3 : p.db1)[wrow0 + threadIdx.x] = colsum;
    stamp(p.dbg, 21);
}

// ============================================================================================================
// post forward:  rows: g1 = LN(g + gh + coef*pool)  |  columns: z4 = g1 W4^T + b4  |  rows: g2 = LN(g1 + gelu(z4))
// ============================================================================================================
struct PostFwd {
    const float *g, *gh, *pool, *Wp, *gamma1, *beta1, *W4, *b4, *gamma2, *beta2;
    float *g1, *xhat1, *rstd1, *z4, *xhat2, *rstd2, *g2;
    int B, G, Kdim;
    unsigned long long* dbg;
};

// LayerNorm of one row held as one value per (thread, i): v[i] is column threadIdx.x + i*512
__device__ __forceinline__ void ln_row(float (&v)[kQ], int G, const float* __restrict__ gamma, const float* __restrict__ beta,
                                       float* xhat_out, float* y_out, float* rstd_out, float* red) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kQ; ++i) s += v[i];
    const float mean = block_sum(s, red) / (float)G;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < kQ; ++i) {
        const int k = threadIdx.x + i * kThreads;
        const float d = k < G ? v[i] - mean : 0.f;
        q += d * d;
    }
    const float rstd = rsqrtf(block_sum(q, red) / (float)G + kEps);
#pragma unroll
    for (int i = 0; i < kQ; ++i) {
        const int k = threadIdx.x + i * kThreads;
        if (k < G) {
            const float xh = (v[i] - mean) * rstd;
            xhat_out[k] = xh;
            y_out[k] = xh * gamma[k] + beta[k];
        }
    }
    if (threadIdx.x == 0) *rstd_out = rstd;
}

__global__ void __launch_bounds__(kThreads, 1) global_post_fwd_kernel(PostFwd p) {
    extern __shared__ __align__(16) float sm[];
    __shared__ float red[32];
    cg::grid_group grid = cg::this_grid();
    const int G = p.G, MT = row_tile_of(G), lda = G + kPad;
    float* A_s = sm;                                   // [MT][lda]  g1 rows
    float* Bt_s = A_s + (size_t)MT * lda;              // [G][8]
    float* red_s = Bt_s + (size_t)G * kNT;             // [512*8]
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    uint32_t phase = 0;
    const int n0 = blockIdx.x * kNT;
    stamp(p.dbg, 48);
    load_bt_rows(Bt_s, p.W4, G, n0, G);
    stamp(p.dbg, 49);
    // ---- row phase: LayerNorm 1
    const float coef = coef_of(p.Wp, p.Kdim, red);
    for (int r = blockIdx.x; r < p.B; r += gridDim.x) {
        const size_t off = (size_t)r * G;
        float v[kQ];
#pragma unroll
        for (int i = 0; i < kQ; ++i) {
            const int k = threadIdx.x + i * kThreads;
            v[i] = k < G ? (p.g[off + k] + p.gh[off + k]) + coef * p.pool[off + k] : 0.f;
        }
        ln_row(v, G, p.gamma1, p.beta1, p.xhat1 + off, p.g1 + off, p.rstd1 + r, red);
    }
    stamp(p.dbg, 50);
    grid.sync();
    stamp(p.dbg, 51);
    // ---- column phase: z4[:, n0..n0+7]
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();
        // g1 was written by other CTAs before the grid barrier (generic proxy); the TMA unit reads it through the async proxy
        if (threadIdx.x < 32) { asm volatile("fence.proxy.async;" ::: "memory"); tma_rows(A_s, lda, p.g1, G, row0, mt, G, &bar); }
        zero_rows(A_s, lda, mt, MT, G);
        mbar_wait(&bar, phase);
        phase ^= 1;
        __syncthreads();
        stamp(p.dbg, 52);
        float out[kNT];
        cta_gemm8(A_s, lda, Bt_s, G, MT, red_s, out);
        stamp(p.dbg, 53);
        if (threadIdx.x < mt) {
            float* o = p.z4 + (size_t)(row0 + threadIdx.x) * G + n0;
            const float4 b0 = ld4(p.b4 + n0), b1 = ld4(p.b4 + n0 + 4);
            *reinterpret_cast<float4*>(o) = make_float4(out[0] + b0.x, out[1] + b0.y, out[2] + b0.z, out[3] + b0.w);
            *reinterpret_cast<float4*>(o + 4) = make_float4(out[4] + b1.x, out[5] + b1.y, out[6] + b1.z, out[7] + b1.w);
        }
    }
    stamp(p.dbg, 54);
    grid.sync();
    stamp(p.dbg, 55);
    // ---- row phase: LayerNorm 2
    for (int r = blockIdx.x; r < p.B; r += gridDim.x) {
        const size_t off = (size_t)r * G;
        float v[kQ];
#pragma unroll
        for (int i = 0; i < kQ; ++i) {
            const int k = threadIdx.x + i * kThreads;
            v[i] = k < G ? __ldcg(p.g1 + off + k) + gelu_f(__ldcg(p.z4 + off + k)) : 0.f;
        }
        ln_row(v, G, p.gamma2, p.beta2, p.xhat2 + off, p.g2 + off, p.rstd2 + r, red);
    }
    stamp(p.dbg, 56);
}

// ============================================================================================================
// post backward:  rows: LN2 backward -> dx2, dz4 = dx2*gelu'(z4)
//                 columns: dg1 = dx2 + dz4 W4 ; dW4 = dz4^T g1 ; db4, dgamma2, dbeta2
//                 rows: LN1 backward -> dx1, dpool, dWp ;  columns (own 8): dgamma1, dbeta1
// ============================================================================================================
struct PostBwd {
    const float *dg2, *gamma2, *xhat2, *rstd2, *z4, *g1, *W4, *gamma1, *xhat1, *rstd1, *pool, *Wp;
    const float* z3;               // optional: pre-activation of gh, to emit dz3 = dx1 * gelu'(z3) for the pre backward kernel
    float* dz3;
    int skip_dw;                   // leave dW4 to the caller (dz4 is scratch3[2], pb_matmul_tn on another stream)
    float *dg1, *dx2, *dz4, *dx1, *dpool, *dWp, *dgamma1, *dbeta1, *dW4, *db4, *dgamma2, *dbeta2;
    int B, G, Kdim;
    unsigned long long* dbg;
};

__global__ void __launch_bounds__(kThreads, 1) global_post_bwd_kernel(PostBwd p) {
    extern __shared__ __align__(16) float sm[];
    __shared__ float red[32];
    cg::grid_group grid = cg::this_grid();
    const int G = p.G, MT = row_tile_of(G), lda = G + kPad;
    float* A_s = sm;                                   // [MT][lda]  dz4 rows
    float* Bt_s = A_s + (size_t)MT * lda;              // [G][8]     W4[:, n0..n0+7]
    float* red_s = Bt_s + (size_t)G * kNT;             // [512*8]
    float* D_s = red_s + kThreads * kNT;               // [MT][8] dz4 slice
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    uint32_t phase = 0;
    const int n0 = blockIdx.x * kNT;
    stamp(p.dbg, 32);
    load_bt_cols(Bt_s, p.W4, G, n0, G);
    if (blockIdx.x == 0) for (int k = threadIdx.x; k < p.Kdim; k += kThreads) p.dWp[k] = 0.f;
    // ---- row phase: LayerNorm-2 backward
    for (int r = blockIdx.x; r < p.B; r += gridDim.x) {
        const size_t off = (size_t)r * G;
        float d[kQ], xh[kQ];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < kQ; ++i) {
            const int k = threadIdx.x + i * kThreads;
            d[i] = k < G ? p.dg2[off + k] * p.gamma2[k] : 0.f;
            xh[i] = k < G ? p.xhat2[off + k] : 0.f;
            s1 += d[i];
            s2 += d[i] * xh[i];
        }
        const float m1 = block_sum(s1, red) / (float)G, m2 = block_sum(s2, red) / (float)G;
        const float rs = p.rstd2[r];
#pragma unroll
        for (int i = 0; i < kQ; ++i) {
            const int k = threadIdx.x + i * kThreads;
            if (k < G) {
                const float dx2 = rs * (d[i] - m1 - xh[i] * m2);
                p.dx2[off + k] = dx2;
                p.dz4[off + k] = dx2 * gelu_grad_f(p.z4[off + k]);
            }
        }
    }
    stamp(p.dbg, 33);
    grid.sync();
    stamp(p.dbg, 34);
    // ---- column phase
    float accw[kQ][kNT];
#pragma unroll
    for (int q = 0; q < kQ; ++q)
#pragma unroll
        for (int j = 0; j < kNT; ++j) accw[q][j] = 0.f;
    float c_db4 = 0.f, c_dg2 = 0.f, c_db2 = 0.f;       // column sums for column n0 + threadIdx.x (< 8)
    for (int row0 = 0; row0 < p.B; row0 += MT) {
        const int mt = min(MT, p.B - row0);
        __syncthreads();
        // dz4 was written by other CTAs before the grid barrier (generic proxy); the TMA unit reads it through the async proxy
        if (threadIdx.x < 32) { asm volatile("fence.proxy.async;" ::: "memory"); tma_rows(A_s, lda, p.dz4, G, row0, mt, G, &bar); }
        zero_rows(A_s, lda, mt, MT, G);
        mbar_wait(&bar, phase);
        phase ^= 1;
        __syncthreads();
        stamp(p.dbg, 35);
        if (!p.skip_dw) fill_x_slice(D_s, p.g1, G, row0, mt, MT, n0);   // g1[:, n0 .. n0+7]
        if (threadIdx.x < 32) {          // one warp: 8 columns x 4 row lanes, folded with two shuffles
            const int j = threadIdx.x & 7, c = n0 + j;
            float a = 0.f, b = 0.f, e = 0.f;
#pragma unroll 4
            for (int m = threadIdx.x >> 3; m < mt; m += 4) {
                const size_t o = (size_t)(row0 + m) * G + c;
                const float dv = p.dg2[o];
                a += A_s[(size_t)m * lda + c];
                b += dv * p.xhat2[o];
                e += dv;
            }
            a += __shfl_xor_sync(0xffffffffu, a, 8); a += __shfl_xor_sync(0xffffffffu, a, 16);
            b += __shfl_xor_sync(0xffffffffu, b, 8); b += __shfl_xor_sync(0xffffffffu, b, 16);
            e += __shfl_xor_sync(0xffffffffu, e, 8); e += __shfl_xor_sync(0xffffffffu, e, 16);
            c_db4 += a; c_dg2 += b; c_db2 += e;       // read by threads < 8 only
        }
        __syncthreads();
        stamp(p.dbg, 36);
        if (!p.skip_dw) dw_from_tile(accw, A_s, lda, D_s, mt, G);       // dW4[:, n0 .. n0+7] = dz4^T g1[:, n0 ..]
        stamp(p.dbg, 37);
        float out[kNT];
        cta_gemm8(A_s, lda, Bt_s, G, MT, red_s, out);                   // dz4 W4[:, n0..]
        stamp(p.dbg, 38);
        if (threadIdx.x < mt) {
            const size_t o = (size_t)(row0 + threadIdx.x) * G + n0;
            const float4 x0 = ld4cg(p.dx2 + o), x1 = ld4cg(p.dx2 + o + 4);
            *reinterpret_cast<float4*>(p.dg1 + o) = make_float4(out[0] + x0.x, out[1] + x0.y, out[2] + x0.z, out[3] + x0.w);
            *reinterpret_cast<float4*>(p.dg1 + o + 4) = make_float4(out[4] + x1.x, out[5] + x1.y, out[6] + x1.z, out[7] + x1.w);
        }
    }
    if (!p.skip_dw) {
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            const int n = threadIdx.x + q * kThreads;
            if (n < G) store8(p.dW4 + (size_t)n * G + n0, accw[q]);
        }
    }
    if (threadIdx.x < kNT) {
        p.db4[n0 + threadIdx.x] = c_db4;
        p.dgamma2[n0 + threadIdx.x] = c_dg2;
        p.dbeta2[n0 + threadIdx.x] = c_db2;
    }
    stamp(p.dbg, 39);
    grid.sync();
    stamp(p.dbg, 40);
    // ---- row phase: LayerNorm-1 backward
    const float coef = coef_of(p.Wp, p.Kdim, red);
    float dcoef = 0.f;
    for (int r = blockIdx.x; r < p.B; r += gridDim.x) {
        const size_t off = (size_t)r * G;
        float d[kQ], xh[kQ];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < kQ; ++i) {
            const int k = threadIdx.x + i * kThreads;
            d[i] = k < G ? __ldcg(p.dg1 + off + k) * p.gamma1[k] : 0.f;
            xh[i] = k < G ? p.xhat1[off + k] : 0.f;
            s1 += d[i];
            s2 += d[i] * xh[i];
        }
        const float m1 = block_sum(s1, red) / (float)G, m2 = block_sum(s2, red) / (float)G;
        const float rs = p.rstd1[r];
#pragma unroll
        for (int i = 0; i < kQ; ++i) {
            const int k = threadIdx.x + i * kThreads;
            if (k < G) {
                const float dx = rs * (d[i] - m1 - xh[i] * m2);
                p.dx1[off + k] = dx;
                if (p.dz3 != nullptr) p.dz3[off + k] = dx * gelu_grad_f(p.z3[off + k]);
                p.dpool[off + k] = coef * dx;
                dcoef += dx * p.pool[off + k];
            }
        }
    }
    const float tot = block_sum(dcoef, red) / (float)p.Kdim;
    if (tot != 0.f)
        for (int k = threadIdx.x; k < p.Kdim; k += kThreads) atomicAdd(&p.dWp[k], tot);
    stamp(p.dbg, 41);
    // ---- own columns: dgamma1 / dbeta1 (needs every row of dg1, complete since the barrier)
    if (threadIdx.x < 32) {
        const int c = n0 + (threadIdx.x & 7);
        float sg = 0.f, sb = 0.f;
#pragma unroll 4
        for (int m = threadIdx.x >> 3; m < p.B; m += 4) {
            const size_t o = (size_t)m * G + c;
            const float dv = __ldcg(p.dg1 + o);
            sg += dv * p.xhat1[o];
            sb += dv;
        }
        sg += __shfl_xor_sync(0xffffffffu, sg, 8); sg += __shfl_xor_sync(0xffffffffu, sg, 16);
        sb += __shfl_xor_sync(0xffffffffu, sb, 8); sb += __shfl_xor_sync(0xffffffffu, sb, 16);
        if (threadIdx.x < kNT) {
            p.dgamma1[c] = sg;
            p.dbeta1[c] = sb;
        }
    }
    stamp(p.dbg, 42);
}

size_t smem_bytes(int MT, int KA, bool with_slice) {
    size_t f = (size_t)MT * (KA + kPad) + (size_t)KA * kNT + (size_t)kThreads * kNT;
    if (with_slice) f += (size_t)MT * kNT;
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
    if (G > kQ * kThreads) return 0;                              // row phases: kQ values per thread
    if ((G + C) / kNT > 148) return 0;                            // cooperative kernels: one resident wave
    if (smem_bytes(row_tile_of(G + C), G + C, true) > 220 * 1024) return 0;
    return 1;
}

int pb_global_pre_fwd(const float* g, const float* W1, const float* b1, const float* W3, const float* b3, float* zg, float* g2l,
                      float* z3, float* gh, int B, int G, int C, void* stream) {
    PB_REQUIRE(pb_global_fused_ok(B, G, C), PB_ERR_UNSUPPORTED, "global_pre_fwd: shape B=%d G=%d C=%d not covered", B, G, C);
    const size_t smem = smem_bytes(row_tile_of(G), G, false);
    int rc = set_smem(global_pre_fwd_kernel, smem);
    if (rc) return rc;
    PreFwd p{g, W1, b1, W3, b3, zg, g2l, z3, gh, B, G, C};
    PB_CUDA(pb_launch_pdl(global_pre_fwd_kernel, dim3((G + C) / kNT), dim3(kThreads), smem, (cudaStream_t)stream, p));
    pb_count_launch(1);
    return pb_check_launch("pb_global_pre_fwd");
}

int pb_global_pre_bwd(const float* dg2l, const float* dgh, const float* zg, const float* z3, const float* g, const float* W1,
                      const float* W3, float* dg, float* dW1, float* db1, float* dW3, float* db3, int B, int G, int C,
                      int dgh_is_dz3, float* dzg_out, void* stream) {
    PB_REQUIRE(pb_global_fused_ok(B, G, C), PB_ERR_UNSUPPORTED, "global_pre_bwd: shape B=%d G=%d C=%d not covered", B, G, C);
    const size_t smem = smem_bytes(row_tile_of(G + C), G + C, true);
    int rc = dgh_is_dz3 ? set_smem(global_pre_bwd_kernel<true>, smem) : set_smem(global_pre_bwd_kernel<false>, smem);
    if (rc) return rc;
    // dW1 == NULL (with dW3 == NULL): the weight gradients are left to the caller, who needs dzg_out for dW1 = dzg^T g
    PB_REQUIRE((dW1 == nullptr) == (dW3 == nullptr), PB_ERR_SHAPE, "global_pre_bwd: dW1 and dW3 go together");
    PB_REQUIRE(dW1 != nullptr || dzg_out != nullptr, PB_ERR_SHAPE, "global_pre_bwd: without dW1/dW3 the caller needs dzg_out");
    PreBwd p{dg2l, dgh, zg, z3, g, W1, W3, dg, dW1, db1, dW3, db3, B, G, C, dgh_is_dz3, dzg_out, dW1 == nullptr ? 1 : 0,
             (unsigned long long*)pb_debug_ptr()};
    if (dgh_is_dz3) PB_CUDA(pb_launch_pdl(global_pre_bwd_kernel<true>, dim3((G + C) / kNT), dim3(kThreads), smem, (cudaStream_t)stream, p));
    else PB_CUDA(pb_launch_pdl(global_pre_bwd_kernel<false>, dim3((G + C) / kNT), dim3(kThreads), smem, (cudaStream_t)stream, p));
    pb_count_launch(1);
    return pb_check_launch("pb_global_pre_bwd");
}

int pb_global_post_fwd(const float* g, const float* gh, const float* pool, const float* Wp, int Kdim, const float* gamma1,
                       const float* beta1, const float* W4, const float* b4, const float* gamma2, const float* beta2, float* g1,
                       float* xhat1, float* rstd1, float* z4, float* xhat2, float* rstd2, float* g2, int B, int G,
                       void* stream) {
    PB_REQUIRE(pb_global_fused_ok(B, G, 8), PB_ERR_UNSUPPORTED, "global_post_fwd: shape B=%d G=%d not covered", B, G);
    PB_REQUIRE(Kdim > 0, PB_ERR_SHAPE, "global_post_fwd: Kdim missing");
    const size_t smem = smem_bytes(row_tile_of(G), G, false);
    int rc = set_smem(global_post_fwd_kernel, smem);
    if (rc) return rc;
    PostFwd p{g, gh, pool, Wp, gamma1, beta1, W4, b4, gamma2, beta2, g1, xhat1, rstd1, z4, xhat2, rstd2, g2, B, G, Kdim,
              (unsigned long long*)pb_debug_ptr()};
    void* args[] = {&p};
    PB_CUDA(cudaLaunchCooperativeKernel((const void*)global_post_fwd_kernel, dim3(G / kNT), dim3(kThreads), args, smem,
                                        (cudaStream_t)stream));
    pb_count_launch(1);
    return pb_check_launch("pb_global_post_fwd");
}

int pb_global_post_bwd(const float* dg2, const float* gamma2, const float* xhat2, const float* rstd2, const float* z4,
                       const float* g1, const float* W4, const float* gamma1, const float* xhat1, const float* rstd1,
                       const float* pool, const float* Wp, int Kdim, float* scratch3, float* dx1, float* dpool, float* dWp,
                       float* dgamma1, float* dbeta1, float* dW4, float* db4, float* dgamma2, float* dbeta2, const float* z3,
                       float* dz3, int B, int G, void* stream) {
    PB_REQUIRE(pb_global_fused_ok(B, G, 8), PB_ERR_UNSUPPORTED, "global_post_bwd: shape B=%d G=%d not covered", B, G);
    PB_REQUIRE(Kdim > 0 && scratch3 != nullptr, PB_ERR_SHAPE, "global_post_bwd: bad arguments");
    const size_t smem = smem_bytes(row_tile_of(G), G, true);
    int rc = set_smem(global_post_bwd_kernel, smem);
    if (rc) return rc;
    const size_t n = (size_t)B * G;
    PB_REQUIRE((z3 == nullptr) == (dz3 == nullptr), PB_ERR_SHAPE, "global_post_bwd: z3 and dz3 go together");
    PostBwd p{dg2, gamma2, xhat2, rstd2, z4, g1, W4, gamma1, xhat1, rstd1, pool, Wp, z3, dz3, dW4 == nullptr ? 1 : 0,
              scratch3, scratch3 + n, scratch3 + 2 * n, dx1, dpool, dWp, dgamma1, dbeta1, dW4, db4, dgamma2, dbeta2, B, G, Kdim, (unsigned long long*)pb_debug_ptr()};
    void* args[] = {&p};
    PB_CUDA(cudaLaunchCooperativeKernel((const void*)global_post_bwd_kernel, dim3(G / kNT), dim3(kThreads), args, smem,
                                        (cudaStream_t)stream));
    pb_count_launch(1);
    return pb_check_launch("pb_global_post_bwd");
}

}  // extern "C"
