// Global track of ProteinBERTBlock as ONE 16-CTA thread-block cluster per block boundary (modules.py:166-199,208,221-229).
//
//   forward unit  F_i = [post of block i ; pre of block i+1]
//        g1 = LN1(g + gh + coef*pool) ; z4 = g1 W4^T + b4 ; g2 = LN2(g1 + gelu(z4))              (:92,:223-229 of block i)
//        [zg | z3] = g2 [W1 ; W3]^T + [b1 ; b3] ; g2l = gelu(zg) ; gh = gelu(z3)                   (:208,:221 of block i+1)
//   backward unit U_i = [pre of block i+1 ; post of block i], the mirror image.
//   Either half may be absent (first block: pre only; last block: post only).
//
// Why a cluster.  The batch is 64 rows, so each product is a [64 x 512] x [512 x 512..640] GEMM: 0.03 - 0.04 GFLOP, pure
// latency.  The cooperative kernels of ops_global_fused.cu spread the output columns over 64-80 CTAs, and every one of
// them pulled the same [64 x 512] A operand out of L2 (2.6 - 2.9 us per staging: 64 requests per cache line queue on
// one L2 slice), separated by 1 - 2 us grid barriers, in four launches per block.  Here
//   * 16 CTAs (4 - 5 column tiles of 8 each) form one cluster: phases are separated by barrier.cluster (hardware,
//     ~0.2 us) instead of a cooperative grid barrier, and no cooperative launch is needed, so the kernel may co-run
//     with the convolutions and the weight-gradient GEMMs;
//   * the A operand is fetched from L2 ONCE per cluster: each CTA issues cp.async.bulk ... .multicast::cluster for the
//     rows it owns and the copies land in all 16 shared memories (every tensor map-free copy here reads this GPU's
//     own memory);
//   * weight slices are private to a CTA and do not depend on the previous kernel: they are requested before
//     griddepcontrol.wait and, for the second product, while the row phase in between runs;
//   * products run on the tensor cores with warp-level mma.sync.m16n8k8 TF32 and the error-compensated three-term split
//     (a = hi + lo: hi*hi + hi*lo + lo*hi), fp32 accumulation: fp32-level accuracy, which the parity mode needs;
//   * the batch is cut into groups of 16 rows, one CLUSTER per group (grid = ceil(B/16) clusters): LayerNorm and the products
//     are row-wise, so the groups never talk to each other -- only the column sums of backward (bias / LayerNorm-affine
//     gradients, dWp) meet, through atomics on outputs the caller zeroes.  Measured on B200: legacy mma.sync TF32 issues
//     at ~14 cycles per SM sub-partition, so one cluster for all 64 rows spent 5.5 us per product; four clusters take
//     a quarter of the rows each;
//   * row phases (LayerNorm forward / backward): CTA `rank` owns row `rank` of its group, one column per thread.
// Weight gradients dW = dZ^T X are not formed here: the caller runs pb_matmul_tn on dz4 / dz3 / dzg (side stream).
#include <cuda.h>
#include "tc_common.cuh"
#include "../../include/pb_b200.h"

void* pb_debug_ptr();          // tc_host.cu: optional device buffer for phase timestamps (probe builds)

namespace {

#ifdef PB_GLOBAL_PROBE
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

constexpr int kCl = 16;            // CTAs per launch = one cluster (non-portable size)
constexpr int kT = 512;            // threads per CTA
constexpr int kMT = 16;            // A tile rows = batch rows per cluster (one m16 row tile)
constexpr int kPadA = 4;           // A / Wt row pitch = K + 4 floats: 16-byte aligned rows, conflict-free fragment reads
constexpr int kMaxNT = 5;          // column tiles (of 8) per CTA
constexpr int kRedPitch = 41;      // partial-tile pitch (floats) in the reduction buffer
constexpr float kEps = 1e-5f;

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
// release global/shared writes to the cluster (and to the async proxy: later TMA reads of what was just written)
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("fence.proxy.async;" ::: "memory");
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// the same copy delivered to the same shared-memory offset (and mbarrier) of every CTA in `mask`; source: this GPU's memory
__device__ __forceinline__ void bulk_g2s_multicast(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
        : "memory");
}

__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
// a = hi + lo with hi = a truncated to TF32 (top 19 bits); lo = a - hi is exact in fp32 and is truncated by the MMA itself
__device__ __forceinline__ void split_tf32(float a, uint32_t& hi, uint32_t& lo) {
    hi = __float_as_uint(a) & 0xffffe000u;
    lo = __float_as_uint(a - __uint_as_float(hi));
}

// columns [c0, c0 + 8*nt) of an N-column product belong to CTA `rank`
__device__ __forceinline__ void my_tiles(int N, int rank, int& c0, int& nt) {
    const int tiles = N >> 3, per = (tiles + kCl - 1) / kCl;
    const int t0 = min(tiles, rank * per), t1 = min(tiles, t0 + per);
    c0 = t0 << 3;
    nt = t1 - t0;
}

struct Smem {
    float* A;        // [kMT][K + 4]   staged A operand (later: the K-group partial tiles)
    float* W;        // forward: [<= 40][K + 4] weight rows ; backward: [K][32] weight columns, 8-float groups XOR-swizzled
    float* red;      // [16][4] row-phase reductions + [512] column-sum scratch
    uint64_t* abar;  // A tile arrival
    uint64_t* wbar;  // weight slice arrival (forward)
};

// ---- A operand: row `rank` of src [Bc, K] (this cluster's rows) -> every CTA of the cluster; rows >= Bc are zero ---------
// Call after a cluster barrier (all CTAs are done with the previous contents); wait with mbar_wait(abar, parity).
__device__ __forceinline__ void stage_a_multicast(const Smem& s, int lda, const float* src, int Bc, int K, int rank) {
    const int kv = K >> 2;
    for (int i = threadIdx.x; i < (kMT - Bc) * kv; i += kT)
        *reinterpret_cast<float4*>(s.A + (size_t)(Bc + i / kv) * lda + (i % kv) * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    if (threadIdx.x == 0) {
        mbar_expect_tx(s.abar, (uint32_t)Bc * (uint32_t)K * 4u);
        if (rank < Bc) bulk_g2s_multicast(s.A + (size_t)rank * lda, src + (size_t)rank * K, (uint32_t)K * 4u, s.abar, 0xffffu);
    }
}

// ---- product: this warp's share of A[16 x K] . B: K group = warp >> 2 (four), column tiles (warp & 3) and (warp & 3) + 4 ---
// kRowsB: B element (k, n) = W[n*ldw + k] (weight rows, forward); else W[k*32 + swizzled(n)] (weight columns, backward).
// Three accumulators per tile (lo*hi, hi*lo, hi*hi) keep the dependent-MMA chains short; they are summed at the end.
template <bool kRowsB>
__device__ __forceinline__ void mma_partial(const float* A, int lda, const float* W, int ldw, int K, int nt, float (&acc)[2][4]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int nq = warp & 3, kg = warp >> 2;
    const int steps = K >> 3, s0 = kg * steps / 4, s1 = (kg + 1) * steps / 4;
    const float* a0p = A + (size_t)g * lda + t;
    const float* a1p = a0p + 8 * lda;
    float c3[2][3][4] = {};
#pragma unroll 2
    for (int st = s0; st < s1; ++st) {
        const int k = st << 3;
        uint32_t ah[4], al[4];
        split_tf32(a0p[k], ah[0], al[0]);
        split_tf32(a1p[k], ah[1], al[1]);
        split_tf32(a0p[k + 4], ah[2], al[2]);
        split_tf32(a1p[k + 4], ah[3], al[3]);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int j = nq + 4 * i;
            if (j < nt) {
                float b0, b1;
                if (kRowsB) {
                    const float* wp = W + (size_t)(j * 8 + g) * ldw + k + t;
                    b0 = wp[0]; b1 = wp[4];
                } else {
                    const int col = ((j ^ t) << 3) + g;          // row k + t (and + 4): k % 8 == 0, so (row & 3) == t for both
                    b0 = W[(size_t)(k + t) * 32 + col];
                    b1 = W[(size_t)(k + t + 4) * 32 + col];
                }
                uint32_t bh[2], bl[2];
                split_tf32(b0, bh[0], bl[0]);
                split_tf32(b1, bh[1], bl[1]);
                mma_tf32(c3[i][0], al, bh);
                mma_tf32(c3[i][1], ah, bl);
                mma_tf32(c3[i][2], ah, bh);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[i][e] += (c3[i][0][e] + c3[i][1][e]) + c3[i][2][e];
}

// the four K groups' partial tiles -> red[kg][row][col] (pitch kRedPitch), aliasing the A buffer: call after __syncthreads
__device__ __forceinline__ void store_partials(float* red, int nt, const float (&acc)[2][4]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int nq = warp & 3, kg = warp >> 2;
    float* base = red + ((size_t)kg * kMT + g) * kRedPitch + 2 * t;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int j = nq + 4 * i;
        if (j < nt) {
            base[j * 8] = acc[i][0]; base[j * 8 + 1] = acc[i][1];
            base[8 * kRedPitch + j * 8] = acc[i][2]; base[8 * kRedPitch + j * 8 + 1] = acc[i][3];
        }
    }
}
__device__ __forceinline__ float sum_partials(const float* red, int r, int c) {
    const float* p = red + (size_t)r * kRedPitch + c;
    return (p[0] + p[(size_t)kMT * kRedPitch]) + (p[(size_t)2 * kMT * kRedPitch] + p[(size_t)3 * kMT * kRedPitch]);
}

// ---- row-phase helper: the whole CTA reduces one value per thread (one batch row, one column per thread) ----------------
__device__ __forceinline__ float cta_sum(float v, float* red /* [16] */) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float r = 0.f;
#pragma unroll
    for (int w = 0; w < kT / 32; ++w) r += red[w];
    return r;
}

// column sums over this cluster's Bc rows for the CTA's nc columns: dst[c] += sum_r val(r, c) (dst zero on entry: the
// clusters of one launch meet here)
template <class F>
__device__ __forceinline__ void column_sum_add(int Bc, int nc, float* dst, F val) {
    if ((int)threadIdx.x < nc) {
        float a = 0.f;
        for (int r = 0; r < Bc; ++r) a += val(r, (int)threadIdx.x);
        atomicAdd(dst + threadIdx.x, a);
    }
}

// the A region also holds the four K groups' partial tiles (4 x 16 x kRedPitch floats), the larger of the two when K is small
__host__ __device__ __forceinline__ size_t a_floats(int K) {
    const size_t a = (size_t)kMT * (K + kPadA), r = (size_t)4 * kMT * kRedPitch;
    return a > r ? a : r;
}
__device__ __forceinline__ Smem carve(float* sm, int K, int wfloats) {
    Smem s;
    s.A = sm;
    s.W = s.A + a_floats(K);
    s.red = s.W + wfloats;
    s.abar = reinterpret_cast<uint64_t*>(s.red + 16 + kT);
    s.wbar = s.abar + 1;
    return s;
}

// =====================================================================================================================
// forward unit
// =====================================================================================================================
struct GtFwd {
    const float *g, *gh, *pool, *Wp;                                  // post half inputs (block i)
    const float *gamma1, *beta1, *W4, *b4, *gamma2, *beta2;
    float *g1, *xhat1, *rstd1, *z4, *xhat2, *rstd2, *g2;              // post half outputs (saved for backward); g2 = g_{i+1}
    const float *g_in, *W1, *b1, *W3, *b3;                            // pre half (block i+1); g_in used when there is no post half
    float *zg, *g2l, *z3, *gh_out;
    int B, G, C, Kdim, has_post, has_pre;
    unsigned long long* dbg;
};

// weight rows [c0, c0 + nc) of the stacked [W1 (C rows) ; W3 (G rows)] matrix (or of W4 when W1 == nullptr) -> s.W rows
__device__ __forceinline__ void load_w_rows(const Smem& s, int ldw, const float* Wa, int rows_a, const float* Wb, int K, int c0, int nc) {
    if (threadIdx.x < 32) {
        if (threadIdx.x == 0) mbar_expect_tx(s.wbar, (uint32_t)nc * (uint32_t)K * 4u);
        __syncwarp();
        for (int j = threadIdx.x; j < nc; j += 32) {
            const int n = c0 + j;
            const float* src = n < rows_a ? Wa + (size_t)n * K : Wb + (size_t)(n - rows_a) * K;
            bulk_g2s(s.W + (size_t)j * ldw, src, (uint32_t)K * 4u, s.wbar);
        }
    }
}

__global__ void __launch_bounds__(kT, 1) gtrack_fwd_kernel(GtFwd p) {
    extern __shared__ __align__(16) float sm[];
    const int G = p.G, C = p.C, lda = G + kPadA;
    const Smem s = carve(sm, G, kMaxNT * 8 * lda);
    const int rank = (int)cluster_rank();
    const int row0 = ((int)blockIdx.x / kCl) * kMT;       // this cluster's first batch row
    const int Bc = min(kMT, p.B - row0);                  // and how many it has
    stamp(p.dbg, 0);
    if (threadIdx.x == 0) { mbar_init(s.abar, 1); mbar_init(s.wbar, 1); fence_barrier_init(); }
    __syncthreads();
    uint32_t apar = 0, wpar = 0;
    int c0, nt;
    // the weight slice of the first product does not depend on the previous kernel
    if (p.has_post) { my_tiles(G, rank, c0, nt); load_w_rows(s, lda, p.W4, G, nullptr, G, c0, nt * 8); }
    else { my_tiles(C + G, rank, c0, nt); load_w_rows(s, lda, p.W1, C, p.W3, G, c0, nt * 8); }
    cluster_sync_all();                                   // every CTA's barriers exist before any multicast copy is issued
    pdl_wait();
    stamp(p.dbg, 1);
    const int k = threadIdx.x;                            // row phases: one column per thread (G <= 512)
    const bool row_ok = rank < Bc, col_ok = row_ok && k < G;
    const size_t roff = (size_t)row0 * G;                 // [B, G] tensors: this cluster's rows start here
    const size_t off = roff + (size_t)rank * G;           // this CTA's row
    const float* gsrc = p.g_in;

    if (p.has_post) {
        // ---- P0 rows: g1 = LN1(g + gh + coef * pool)
        {
            float tsum = 0.f;
            for (int q = threadIdx.x; q < p.Kdim; q += kT) tsum += p.Wp[q];
            const float coef = cta_sum(tsum, s.red) / (float)p.Kdim;
            const float v = col_ok ? (p.g[off + k] + p.gh[off + k]) + coef * p.pool[off + k] : 0.f;
            const float mean = cta_sum(v, s.red) / (float)G;
            const float d = (k < G) ? v - mean : 0.f;
            const float rstd = rsqrtf(cta_sum(d * d, s.red) / (float)G + kEps);
            if (col_ok) {
                const float xh = d * rstd;
                p.xhat1[off + k] = xh;
                p.g1[off + k] = xh * p.gamma1[k] + p.beta1[k];
            }
            if (row_ok && k == 0) p.rstd1[row0 + rank] = rstd;
        }
        stamp(p.dbg, 2);
        cluster_sync_all();
        stamp(p.dbg, 3);
        // ---- P1 columns: z4 = g1 W4^T + b4 ; v2 = g1 + gelu(z4) (parked in the g2 buffer)
        stage_a_multicast(s, lda, p.g1 + roff, Bc, G, rank);
        mbar_wait(s.wbar, wpar); wpar ^= 1;
        stamp(p.dbg, 4);
        mbar_wait(s.abar, apar); apar ^= 1;
        __syncthreads();
        stamp(p.dbg, 5);
        float acc[2][4] = {};
        mma_partial<true>(s.A, lda, s.W, lda, G, nt, acc);
        __syncthreads();                                  // A and W are free
        stamp(p.dbg, 6);
        if (p.has_pre) {                                  // next product's weights arrive while the row phase runs
            int c0n, ntn;
            my_tiles(C + G, rank, c0n, ntn);
            load_w_rows(s, lda, p.W1, C, p.W3, G, c0n, ntn * 8);
        }
        store_partials(s.A, nt, acc);
        __syncthreads();
        {
            const int nc = nt * 8;
            for (int idx = threadIdx.x; idx < Bc * nc; idx += kT) {
                const int r = idx / nc, c = idx - r * nc;
                const size_t o = roff + (size_t)r * G + c0 + c;
                const float z = sum_partials(s.A, r, c) + p.b4[c0 + c];
                p.z4[o] = z;
                p.g2[o] = __ldcg(p.g1 + o) + gelu_f(z);   // g1 rows were written before the cluster barrier
            }
        }
        stamp(p.dbg, 7);
        cluster_sync_all();
        stamp(p.dbg, 8);
        // ---- P2 rows: g2 = LN2(v2)
        {
            const float v = col_ok ? __ldcg(p.g2 + off + k) : 0.f;
            const float mean = cta_sum(v, s.red) / (float)G;
            const float d = (k < G) ? v - mean : 0.f;
            const float rstd = rsqrtf(cta_sum(d * d, s.red) / (float)G + kEps);
            if (col_ok) {
                const float xh = d * rstd;
                p.xhat2[off + k] = xh;
                p.g2[off + k] = xh * p.gamma2[k] + p.beta2[k];
            }
            if (row_ok && k == 0) p.rstd2[row0 + rank] = rstd;
        }
        gsrc = p.g2;
        stamp(p.dbg, 9);
        if (p.has_pre) cluster_sync_all();
        stamp(p.dbg, 10);
    }

    if (p.has_pre) {
        // ---- P3 columns: [zg | z3] = g [W1 ; W3]^T + [b1 ; b3]
        my_tiles(C + G, rank, c0, nt);
        stage_a_multicast(s, lda, gsrc + roff, Bc, G, rank);
        mbar_wait(s.wbar, wpar); wpar ^= 1;
        mbar_wait(s.abar, apar); apar ^= 1;
        __syncthreads();
        stamp(p.dbg, 11);
        float acc[2][4] = {};
        mma_partial<true>(s.A, lda, s.W, lda, G, nt, acc);
        __syncthreads();
        stamp(p.dbg, 12);
        store_partials(s.A, nt, acc);
        __syncthreads();
        const int nc = nt * 8;
        for (int idx = threadIdx.x; idx < Bc * nc; idx += kT) {
            const int r = idx / nc, c = idx - r * nc, n = c0 + c;
            const float acc_v = sum_partials(s.A, r, c);
            if (n < C) {
                const float z = acc_v + p.b1[n];
                p.zg[(size_t)(row0 + r) * C + n] = z;
                p.g2l[(size_t)(row0 + r) * C + n] = gelu_f(z);
            } else {
                const float z = acc_v + p.b3[n - C];
                p.z3[roff + (size_t)r * G + n - C] = z;
                p.gh_out[roff + (size_t)r * G + n - C] = gelu_f(z);
            }
        }
        stamp(p.dbg, 13);
    }
    cluster_sync_all();          // no CTA leaves while copies it issued, or copies aimed at it, may be in flight
}

// =====================================================================================================================
// backward unit
// =====================================================================================================================
struct GtBwd {
    // pre half (block i+1): dg_total = dg_direct + dz3 W3 + dzg W1, dzg = dg2l * gelu'(zg)
    const float *dg2l, *zg, *dz3_in, *dgh, *z3n, *dg_direct, *W1, *W3;   // dz3_in ready, or dgh * gelu'(z3n) is formed here
    float *dz3_buf, *dzg, *db1, *db3, *dg_total;
    // post half (block i)
    const float *dg_in;                                               // gradient of g2 when there is no pre half
    const float *gamma2, *xhat2, *rstd2, *z4, *W4, *gamma1, *xhat1, *rstd1, *pool, *Wp, *z3;
    float *dv2, *dz4, *dg1, *dv1, *dz3_out, *dpool, *dWp, *dgamma1, *dbeta1, *db4, *dgamma2, *dbeta2;
    int B, G, C, Kdim, has_pre, has_post;
    unsigned long long* dbg;
};

// weight columns [c0, c0 + 8*nt) of W [rows x ldw] -> Ws[(k0 + k)][32], 8-float groups XOR-swizzled by (row & 3)
__device__ __forceinline__ void load_w_cols(float* Ws, int k0, const float* __restrict__ W, int rows, int ldw, int c0, int nt) {
    const int per_row = nt * 2;                                       // float4 per row
    for (int i = threadIdx.x; i < rows * per_row; i += kT) {
        const int k = i / per_row, q = i - k * per_row;               // q: float4 index inside the row's slice
        const float4 v = __ldg(reinterpret_cast<const float4*>(W + (size_t)k * ldw + c0 + q * 4));
        const int row = k0 + k;
        const int col = (((q >> 1) ^ (row & 3)) << 3) + (q & 1) * 4;
        *reinterpret_cast<float4*>(Ws + (size_t)row * 32 + col) = v;
    }
}

__global__ void __launch_bounds__(kT, 1) gtrack_bwd_kernel(GtBwd p) {
    extern __shared__ __align__(16) float sm[];
    const int G = p.G, C = p.C, lda = G + kPadA;
    const Smem s = carve(sm, G, (G + C) * 32);
    const int rank = (int)cluster_rank();
    const int row0 = ((int)blockIdx.x / kCl) * kMT;
    const int Bc = min(kMT, p.B - row0);
    if (threadIdx.x == 0) { mbar_init(s.abar, 1); mbar_init(s.wbar, 1); fence_barrier_init(); }
    __syncthreads();
    uint32_t apar = 0;
    int c0, nt;
    my_tiles(G, rank, c0, nt);                            // every product of the backward unit has G output columns
    const int nc = nt * 8;
    if (p.has_pre) { load_w_cols(s.W, 0, p.W3, G, G, c0, nt); load_w_cols(s.W, G, p.W1, C, G, c0, nt); }
    else load_w_cols(s.W, 0, p.W4, G, G, c0, nt);
    cluster_sync_all();
    pdl_wait();
    const int k = threadIdx.x;
    const bool row_ok = rank < Bc, col_ok = row_ok && k < G;
    const size_t roff = (size_t)row0 * G, off = roff + (size_t)rank * G;
    const float* dg2 = p.dg_in;

    if (p.has_pre) {
        const float* dz3 = p.dz3_in;
        if (dz3 == nullptr) {                             // rows: dz3 = dgh * gelu'(z3) of block i+1
            if (col_ok) p.dz3_buf[off + k] = p.dgh[off + k] * gelu_grad_f(p.z3n[off + k]);
            dz3 = p.dz3_buf;
            cluster_sync_all();
        }
        // ---- Q0 columns: dg_total = dg_direct + dz3 W3 + dzg W1
        stage_a_multicast(s, lda, dz3 + roff, Bc, G, rank);
        mbar_wait(s.abar, apar); apar ^= 1;
        __syncthreads();
        float acc[2][4] = {};
        mma_partial<false>(s.A, lda, s.W, 32, G, nt, acc);
        column_sum_add(Bc, nc, p.db3 + c0, [&](int r, int c) { return s.A[(size_t)r * lda + c0 + c]; });
        __syncthreads();
        // second pass, K = C: A = dzg = dg2l * gelu'(zg), formed by every CTA (16 x C elements)
        const int ldc = C + kPadA;
        for (int i = threadIdx.x; i < kMT * C; i += kT) {
            const int r = i / C, c = i - r * C;
            float v = 0.f;
            if (r < Bc) v = p.dg2l[(size_t)(row0 + r) * C + c] * gelu_grad_f(p.zg[(size_t)(row0 + r) * C + c]);
            s.A[(size_t)r * ldc + c] = v;
        }
        __syncthreads();
        mma_partial<false>(s.A, ldc, s.W + (size_t)G * 32, 32, C, nt, acc);
        {   // db1 / dzg: the columns of dzg are split over the CTAs by their own tiling of C
            int cc0, cnt;
            my_tiles(C, rank, cc0, cnt);
            const int cnc = cnt * 8;
            column_sum_add(Bc, cnc, p.db1 + cc0, [&](int r, int c) { return s.A[(size_t)r * ldc + cc0 + c]; });
            for (int idx = threadIdx.x; idx < Bc * cnc; idx += kT) {
                const int r = idx / cnc, c = idx - r * cnc;
                p.dzg[(size_t)(row0 + r) * C + cc0 + c] = s.A[(size_t)r * ldc + cc0 + c];
            }
        }
        __syncthreads();
        if (p.has_post) load_w_cols(s.W, 0, p.W4, G, G, c0, nt);      // W3/W1 slices are dead: bring in W4's
        store_partials(s.A, nt, acc);
        __syncthreads();
        for (int idx = threadIdx.x; idx < Bc * nc; idx += kT) {
            const int r = idx / nc, c = idx - r * nc;
            const size_t o = roff + (size_t)r * G + c0 + c;
            p.dg_total[o] = sum_partials(s.A, r, c) + (p.dg_direct != nullptr ? p.dg_direct[o] : 0.f);
        }
        dg2 = p.dg_total;
        if (p.has_post) cluster_sync_all();
    }

    if (p.has_post) {
        // ---- Q1 rows: LayerNorm-2 backward -> dv2 ; dz4 = dv2 * gelu'(z4)
        {
            const float d = col_ok ? __ldcg(dg2 + off + k) * p.gamma2[k] : 0.f;
            const float xh = col_ok ? p.xhat2[off + k] : 0.f;
            const float m1 = cta_sum(d, s.red) / (float)G, m2 = cta_sum(d * xh, s.red) / (float)G;
            if (col_ok) {
                const float dv = p.rstd2[row0 + rank] * (d - m1 - xh * m2);
                p.dv2[off + k] = dv;
                p.dz4[off + k] = dv * gelu_grad_f(p.z4[off + k]);
            }
        }
        cluster_sync_all();
        // ---- Q2 columns: dg1 = dv2 + dz4 W4 ; db4, dgamma2, dbeta2
        stage_a_multicast(s, lda, p.dz4 + roff, Bc, G, rank);
        column_sum_add(Bc, nc, p.dgamma2 + c0, [&](int r, int c) {
            const size_t o = roff + (size_t)r * G + c0 + c;
            return __ldcg(dg2 + o) * p.xhat2[o];
        });
        column_sum_add(Bc, nc, p.dbeta2 + c0, [&](int r, int c) { return __ldcg(dg2 + roff + (size_t)r * G + c0 + c); });
        mbar_wait(s.abar, apar); apar ^= 1;
        __syncthreads();
        float acc[2][4] = {};
        mma_partial<false>(s.A, lda, s.W, 32, G, nt, acc);
        column_sum_add(Bc, nc, p.db4 + c0, [&](int r, int c) { return s.A[(size_t)r * lda + c0 + c]; });
        __syncthreads();
        store_partials(s.A, nt, acc);
        __syncthreads();
        for (int idx = threadIdx.x; idx < Bc * nc; idx += kT) {
            const int r = idx / nc, c = idx - r * nc;
            const size_t o = roff + (size_t)r * G + c0 + c;
            p.dg1[o] = sum_partials(s.A, r, c) + __ldcg(p.dv2 + o);
        }
        cluster_sync_all();
        // ---- Q3 rows: LayerNorm-1 backward -> dv1 (gradient of g, of gh and, times coef, of pool), dz3 of THIS block, dWp
        {
            float tsum = 0.f;
            for (int q = threadIdx.x; q < p.Kdim; q += kT) tsum += p.Wp[q];
            const float coef = cta_sum(tsum, s.red) / (float)p.Kdim;
            const float d = col_ok ? __ldcg(p.dg1 + off + k) * p.gamma1[k] : 0.f;
            const float xh = col_ok ? p.xhat1[off + k] : 0.f;
            const float m1 = cta_sum(d, s.red) / (float)G, m2 = cta_sum(d * xh, s.red) / (float)G;
            float dcoef = 0.f;
            if (col_ok) {
                const float dv = p.rstd1[row0 + rank] * (d - m1 - xh * m2);
                p.dv1[off + k] = dv;
                if (p.dz3_out != nullptr) p.dz3_out[off + k] = dv * gelu_grad_f(p.z3[off + k]);
                p.dpool[off + k] = coef * dv;
                dcoef = dv * p.pool[off + k];
            }
            const float tot = cta_sum(dcoef, s.red) / (float)p.Kdim;
            if (tot != 0.f)
                for (int q = threadIdx.x; q < p.Kdim; q += kT) atomicAdd(&p.dWp[q], tot);       // dWp zero on entry
        }
        // own columns: dgamma1 / dbeta1 need this cluster's rows of dg1 (complete since the barrier before Q3)
        column_sum_add(Bc, nc, p.dgamma1 + c0, [&](int r, int c) {
            const size_t o = roff + (size_t)r * G + c0 + c;
            return __ldcg(p.dg1 + o) * p.xhat1[o];
        });
        column_sum_add(Bc, nc, p.dbeta1 + c0, [&](int r, int c) { return __ldcg(p.dg1 + roff + (size_t)r * G + c0 + c); });
    }
    cluster_sync_all();
}

size_t fwd_smem(int G) { return (a_floats(G) + (size_t)kMaxNT * 8 * (G + kPadA) + 16 + kT) * sizeof(float) + 32; }
size_t bwd_smem(int G, int C) { return (a_floats(G) + (size_t)(G + C) * 32 + 16 + kT) * sizeof(float) + 32; }

template <class Kern, class P>
int launch_cluster(Kern kern, const P& p, size_t smem, cudaStream_t st, const char* what) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kCl * ((p.B + kMT - 1) / kMT)); cfg.blockDim = dim3(kT); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kCl; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pb_pdl_enabled() ? 2 : 1;
    PB_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
    pb_count_launch(1);
    return pb_check_launch(what);
}

// one-time per kernel: opt in to the shared memory and the 16-CTA cluster, and ask the driver whether such a cluster fits
template <class Kern>
int prepare(Kern kern, size_t smem, int* state /* 0 unknown, 1 ok, -1 unavailable */, size_t* smem_set) {
    if (*state == 0 || *smem_set < smem) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
            cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
            cudaGetLastError();
            *state = -1;
            return *state;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(kCl); cfg.blockDim = dim3(kT); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = kCl; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { cudaGetLastError(); n = 0; }
        *state = n >= 1 ? 1 : -1;
        *smem_set = smem;
    }
    return *state;
}

int g_fwd_state = 0, g_bwd_state = 0;
size_t g_fwd_smem = 0, g_bwd_smem = 0;


// ---- weight gradients of the global track: up to three products out_j[n, k] = sum_{m < M} A_j[m, n] X_j[m, k] per launch ---
// M = batch (<= 128) is the contraction, so each 64 x 64 output tile needs only A[:, n-tile] and X[:, k-tile] (M x 64 each):
// one CTA per tile, 4 x 4 outputs per thread, blockIdx.z picks the product.  Replaces three pb_matmul_tn launches per unit.
struct DwJob { const float* A; const float* X; float* out; int N, K; };
struct DwJobs { DwJob j[3]; int M; };

__global__ void __launch_bounds__(256) gtrack_dw_kernel(DwJobs p) {
    __shared__ float As[64 * 64], Xs[64 * 64];
    const DwJob jb = p.j[blockIdx.z];
    const int n0 = blockIdx.x * 64, k0 = blockIdx.y * 64;
    if (n0 >= jb.N || k0 >= jb.K) return;
    pdl_wait();
    const int tn = (threadIdx.x >> 4) * 4, tk = (threadIdx.x & 15) * 4;
    float acc[4][4] = {};
    for (int m0 = 0; m0 < p.M; m0 += 64) {            // the batch in chunks of 64 rows
        const int mc = min(64, p.M - m0);
        __syncthreads();
        for (int i = threadIdx.x; i < mc * 16; i += 256) {
            const int m = i >> 4, q = (i & 15) * 4;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), x = a;
            if (n0 + q < jb.N) a = *reinterpret_cast<const float4*>(jb.A + (size_t)(m0 + m) * jb.N + n0 + q);       // N, K multiples of 4
            if (k0 + q < jb.K) x = *reinterpret_cast<const float4*>(jb.X + (size_t)(m0 + m) * jb.K + k0 + q);
            *reinterpret_cast<float4*>(As + m * 64 + q) = a;
            *reinterpret_cast<float4*>(Xs + m * 64 + q) = x;
        }
        __syncthreads();
#pragma unroll 4
        for (int m = 0; m < mc; ++m) {
            const float4 a = *reinterpret_cast<const float4*>(As + m * 64 + tn);
            const float4 x = *reinterpret_cast<const float4*>(Xs + m * 64 + tk);
            const float av[4] = {a.x, a.y, a.z, a.w}, xv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], xv[j], acc[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int n = n0 + tn + i;
        if (n < jb.N && k0 + tk < jb.K)
            *reinterpret_cast<float4*>(jb.out + (size_t)n * jb.K + k0 + tk) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
}

}  // namespace

extern "C" {

int pb_matmul_tn3(const float* A0, const float* X0, float* out0, int N0, int K0, const float* A1, const float* X1, float* out1,
                  int N1, int K1, const float* A2, const float* X2, float* out2, int N2, int K2, int M, void* stream) {
    DwJobs p{};
    p.M = M;
    const float* As[3] = {A0, A1, A2};
    const float* Xs[3] = {X0, X1, X2};
    float* outs[3] = {out0, out1, out2};
    const int Ns[3] = {N0, N1, N2}, Ks[3] = {K0, K1, K2};
    int n = 0, maxN = 0, maxK = 0;
    for (int i = 0; i < 3; ++i) {
        if (outs[i] == nullptr) continue;
        PB_REQUIRE(As[i] && Xs[i] && Ns[i] > 0 && Ks[i] > 0 && Ns[i] % 4 == 0 && Ks[i] % 4 == 0, PB_ERR_SHAPE,
                   "matmul_tn3: product %d needs A, X and dimensions that are multiples of 4", i);
        p.j[n++] = DwJob{As[i], Xs[i], outs[i], Ns[i], Ks[i]};
        maxN = Ns[i] > maxN ? Ns[i] : maxN;
        maxK = Ks[i] > maxK ? Ks[i] : maxK;
    }
    if (n == 0) return PB_OK;
    PB_REQUIRE(M >= 1 && M <= 128, PB_ERR_SHAPE, "matmul_tn3: contraction length M = %d outside 1..128", M);
    PB_CUDA(pb_launch_pdl(gtrack_dw_kernel, dim3((maxN + 63) / 64, (maxK + 63) / 64, n), dim3(256), 0, (cudaStream_t)stream, p));
    pb_count_launch(1);
    return pb_check_launch("pb_matmul_tn3");
}

int pb_gtrack_ok(int B, int G, int C) {
    if (B < 1 || B > 128 || G < 8 || C < 8 || G % 8 || C % 8 || G > 512 || C > G) return 0;
    if (((C + G) / 8 + kCl - 1) / kCl > kMaxNT) return 0;
    if (fwd_smem(G) > 232448 || bwd_smem(G, C) > 232448) return 0;
    if (prepare(gtrack_fwd_kernel, fwd_smem(G), &g_fwd_state, &g_fwd_smem) != 1) return 0;
    if (prepare(gtrack_bwd_kernel, bwd_smem(G, C), &g_bwd_state, &g_bwd_smem) != 1) return 0;
    return 1;
}

int pb_gtrack_fwd(const float* g, const float* gh, const float* pool, const float* Wp, int Kdim, const float* gamma1,
                  const float* beta1, const float* W4, const float* b4, const float* gamma2, const float* beta2, float* g1,
                  float* xhat1, float* rstd1, float* z4, float* xhat2, float* rstd2, float* g2, const float* g_in,
                  const float* W1, const float* b1, const float* W3, const float* b3, float* zg, float* g2l, float* z3,
                  float* gh_out, int B, int G, int C, void* stream) {
    PB_REQUIRE(pb_gtrack_ok(B, G, C), PB_ERR_UNSUPPORTED, "gtrack_fwd: shape B=%d G=%d C=%d not covered (or no 16-CTA cluster)", B, G, C);
    GtFwd p{};
    p.has_post = W4 != nullptr; p.has_pre = W1 != nullptr;
    PB_REQUIRE(p.has_post || p.has_pre, PB_ERR_SHAPE, "gtrack_fwd: nothing to do");
    PB_REQUIRE(!p.has_post || (g && gh && pool && Wp && Kdim > 0 && gamma1 && beta1 && b4 && gamma2 && beta2 && g1 && xhat1 && rstd1 && z4 &&
                               xhat2 && rstd2 && g2), PB_ERR_SHAPE, "gtrack_fwd: post half needs all of its tensors");
    PB_REQUIRE(!p.has_pre || (b1 && W3 && b3 && zg && g2l && z3 && gh_out && (p.has_post || g_in)), PB_ERR_SHAPE,
               "gtrack_fwd: pre half needs all of its tensors");
    p.g = g; p.gh = gh; p.pool = pool; p.Wp = Wp; p.Kdim = Kdim; p.gamma1 = gamma1; p.beta1 = beta1; p.W4 = W4; p.b4 = b4;
    p.gamma2 = gamma2; p.beta2 = beta2; p.g1 = g1; p.xhat1 = xhat1; p.rstd1 = rstd1; p.z4 = z4; p.xhat2 = xhat2; p.rstd2 = rstd2;
    p.g2 = g2; p.g_in = g_in; p.W1 = W1; p.b1 = b1; p.W3 = W3; p.b3 = b3; p.zg = zg; p.g2l = g2l; p.z3 = z3; p.gh_out = gh_out;
    p.B = B; p.G = G; p.C = C;
    p.dbg = (unsigned long long*)pb_debug_ptr();
    return launch_cluster(gtrack_fwd_kernel, p, fwd_smem(G), (cudaStream_t)stream, "pb_gtrack_fwd");
}

int pb_gtrack_bwd(const float* dg2l, const float* zg, const float* dz3_in, const float* dgh, const float* z3n,
                  const float* dg_direct, const float* W1, const float* W3, float* dz3_buf, float* dzg, float* db1, float* db3,
                  float* dg_total, const float* dg_in, const float* gamma2, const float* xhat2, const float* rstd2,
                  const float* z4, const float* W4, const float* gamma1, const float* xhat1, const float* rstd1,
                  const float* pool, const float* Wp, int Kdim, const float* z3, float* dv2, float* dz4, float* dg1, float* dv1,
                  float* dz3_out, float* dpool, float* dWp, float* dgamma1, float* dbeta1, float* db4, float* dgamma2,
                  float* dbeta2, int B, int G, int C, void* stream) {
    PB_REQUIRE(pb_gtrack_ok(B, G, C), PB_ERR_UNSUPPORTED, "gtrack_bwd: shape B=%d G=%d C=%d not covered (or no 16-CTA cluster)", B, G, C);
    GtBwd p{};
    p.has_pre = W1 != nullptr; p.has_post = W4 != nullptr;
    PB_REQUIRE(p.has_post || p.has_pre, PB_ERR_SHAPE, "gtrack_bwd: nothing to do");
    PB_REQUIRE(!p.has_pre || (dg2l && zg && (dz3_in || (dgh && z3n && dz3_buf)) && W3 && dzg && db1 && db3 && dg_total),
               PB_ERR_SHAPE, "gtrack_bwd: pre half needs all of its tensors");
    PB_REQUIRE(!p.has_post || ((p.has_pre || dg_in) && gamma2 && xhat2 && rstd2 && z4 && gamma1 && xhat1 && rstd1 && pool && Wp && Kdim > 0 &&
                               dv2 && dz4 && dg1 && dv1 && dpool && dWp && dgamma1 && dbeta1 && db4 && dgamma2 && dbeta2 &&
                               ((z3 == nullptr) == (dz3_out == nullptr))), PB_ERR_SHAPE, "gtrack_bwd: post half needs all of its tensors");
    p.dg2l = dg2l; p.zg = zg; p.dz3_in = dz3_in; p.dgh = dgh; p.z3n = z3n; p.dg_direct = dg_direct; p.W1 = W1; p.W3 = W3;
    p.dz3_buf = dz3_buf; p.dzg = dzg; p.db1 = db1; p.db3 = db3; p.dg_total = dg_total; p.dg_in = dg_in; p.gamma2 = gamma2;
    p.xhat2 = xhat2; p.rstd2 = rstd2; p.z4 = z4; p.W4 = W4; p.gamma1 = gamma1; p.xhat1 = xhat1; p.rstd1 = rstd1; p.pool = pool;
    p.Wp = Wp; p.Kdim = Kdim; p.z3 = z3; p.dv2 = dv2; p.dz4 = dz4; p.dg1 = dg1; p.dv1 = dv1; p.dz3_out = dz3_out; p.dpool = dpool;
    p.dWp = dWp; p.dgamma1 = dgamma1; p.dbeta1 = dbeta1; p.db4 = db4; p.dgamma2 = dgamma2; p.dbeta2 = dbeta2;
    p.B = B; p.G = G; p.C = C;
    p.dbg = (unsigned long long*)pb_debug_ptr();
    return launch_cluster(gtrack_bwd_kernel, p, bwd_smem(G, C), (cudaStream_t)stream, "pb_gtrack_bwd");
}

}  // extern "C"
