// bf16 tcgen05 GEMMs of the local track that are not convolutions.
//
//   tc_gemm_kernel  : D[128 x 128 tile] = A[M,K] * B[N,K]^T, both operands K-major, fused epilogues:
//       LINEAR_FWD  z = D + b ; u2 = h + gelu(z) ; LayerNorm sums           (modules.py:214-216)
//       POOL_FWD    pool[b, n] += sum over the tile's rows of gelu(D)        (modules.py:55-60,92)
//       POOL_DZ     dzv = dpool[b, n] * gelu'(D)
//       ADD_STORE   out = base + D                                           (data gradients)
//   tc_wgrad_kernel : D[m, n] += sum_rows P[row, m] * Q[row + shift, n]  -- the contraction runs over sequence
//       positions, so both operands are MN-major in shared memory (rows of the activation tensors are K);
//       used for the conv weight gradients (one launch covers all taps), dW of the local linear and dWv.
//
// Same skeleton as tc_conv.cu: warp 0 TMA producer, warp 1 TMEM alloc + MMA issue, warps 2..5 epilogue.
#include "tc_common.cuh"
#include "tc_api.cuh"

namespace {

constexpr int kBM = 128, kBN = 128;
constexpr int kStages = 3;
constexpr int kStageBytes = (kBM + kBN) * 64 * 2;          // 32 KB
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256 + 4 * kBN * 4;
constexpr int kThreads = 192;

enum { EPI_LINEAR_FWD = 0, EPI_POOL_FWD = 1, EPI_POOL_DZ = 2, EPI_ADD_STORE = 3, EPI_STORE_F32_32 = 4, EPI_STORE = 5 };

struct GemmEpi {
    int ldn;                       // row pitch (elements) of the [M, ldn] tensors touched by the epilogue
    int L;                         // rows per sample (sample index = row / L)
    const float* bias;             // [N]
    const bf16* base;              // residual / addend [M, ldn]
    bf16* out0;                    // z / dzv / out
    bf16* out1;                    // u2
    double* stats;                 // [B, 2]
    float* pool;                   // [B, ldn]
    const float* dpool;            // [B, ldn]
    float* f32out;                 // EPI_STORE_F32_32: [M, 32] fp32 (first 32 columns of the tile)
    int base_in_smem;              // LINEAR_FWD with K == N == 128: the residual h tile is the A operand still in the ring
};

// The two pooling GEMMs (K = C = 128: two 64-wide chunks) are bound by their GELU epilogues, i.e. by how many
// epilogue warps an SM holds: they take a two-stage ring so that three CTAs fit per SM.
template <int EPI> struct GemmCfg {
    static constexpr int kStagesE = (EPI == EPI_POOL_FWD || EPI == EPI_POOL_DZ) ? 2 : 3;
    static constexpr int kCtas = (EPI == EPI_POOL_FWD || EPI == EPI_POOL_DZ) ? 3 : 2;
    static constexpr int kSmem = kStagesE * kStageBytes + 1024 + 256 + 4 * kBN * 4;
};

template <int EPI>
__global__ void __launch_bounds__(kThreads, GemmCfg<EPI>::kCtas)
tc_gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
               const __grid_constant__ CUtensorMap mapS0, const __grid_constant__ CUtensorMap mapS1, GemmEpi ep, int K) {
    constexpr int kStages = GemmCfg<EPI>::kStagesE;          // shadows the file-level ring depth
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + kStages * kStageBytes);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 1);
    float* colsum = (float*)(smem + kStages * kStageBytes + 256);      // [4][kBN]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * kBM, n0 = blockIdx.y * kBN;
    const int n_iter = K / 64;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA); tma_prefetch_desc(&mapB);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kBN>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            for (int i = 0; i < n_iter; ++i) {
                const int s = i % kStages;
                mbar_wait(&empty_bar[s], ((i / kStages) & 1) ^ 1);
                uint8_t* a_dst = smem + s * kStageBytes;
                uint8_t* b_dst = a_dst + kBM * 64 * 2;
                mbar_expect_tx(&full_bar[s], kStageBytes);
                tma_load_2d(a_dst, &mapA, &full_bar[s], i * 64, m0);
                tma_load_2d(b_dst, &mapB, &full_bar[s], i * 64, n0);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(kBM, kBN, 0, 0);
            for (int i = 0; i < n_iter; ++i) {
                const int s = i % kStages;
                mbar_wait(&full_bar[s], (i / kStages) & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
                const uint64_t da0 = make_smem_desc_sw128(a_addr, 16, 1024);
                const uint64_t db0 = make_smem_desc_sw128(a_addr + kBM * 64 * 2, 16, 1024);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(tmem_base, da0 + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), idesc, (i == 0 && k == 0) ? 0u : 1u);
                umma_commit(&empty_bar[s]);
            }
            umma_commit(tmem_full_bar);
        }
    } else {
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        const int q = warp & 3;
        const int row = m0 + q * 32 + lane;
        const int b = m0 / ep.L;                         // tiles never straddle samples (L % 128 == 0)
        const size_t grow = (size_t)row * ep.ldn + n0;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        // bf16 outputs leave through TMA: each warp stages its 32 x 32 sub-tile (64-byte swizzle) in the last ring
        // stage, idle by now, and one lane issues the tensor store; two buffers so a store drains behind the math
        constexpr bool kTmaOut = EPI == EPI_LINEAR_FWD || EPI == EPI_POOL_DZ || EPI == EPI_STORE || EPI == EPI_ADD_STORE;
        uint8_t* stage = smem + (kStages - 1) * kStageBytes + (warp - 2) * 8192;
        const int sw = (lane >> 1) & 3;
        const int trow = q * 32 + lane;                              // row inside the tile
        float s_sum = 0.f, s_sq = 0.f;
#pragma unroll 1
        for (int ch = 0; ch < kBN / 32; ++ch) {
            float acc[32];
            tmem_ld32(taddr + ch * 32, acc);
            const int c0 = ch * 32;
            uint8_t* sbuf = stage + (ch & 1) * 4096;
            if (kTmaOut && ch >= 2) {
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
            }
            if (EPI == EPI_LINEAR_FWD) {
                uint4 hv[4];
                if (ep.base_in_smem) {
                    const uint8_t* hrow = smem + (ch >> 1) * kStageBytes + (size_t)trow * 128;
#pragma unroll
                    for (int v = 0; v < 4; ++v) hv[v] = *reinterpret_cast<const uint4*>(hrow + ((((ch & 1) * 4 + v) ^ (trow & 7)) * 16));
                } else {
#pragma unroll
                    for (int v = 0; v < 4; ++v) hv[v] = *reinterpret_cast<const uint4*>(ep.base + grow + c0 + v * 8);
                }
                const bf16* hb = reinterpret_cast<const bf16*>(hv);
                const float bl = __ldg(ep.bias + n0 + c0 + lane);
                __align__(16) bf16 oz[32], ou[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const bf16 z = __float2bfloat16_rn(acc[j] + __shfl_sync(0xffffffffu, bl, j));
                    oz[j] = z;
                    acc[j] = __bfloat162float(z);
                }
                gelu_fast_n(acc);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const bf16 ub = __float2bfloat16_rn(__bfloat162float(hb[j]) + acc[j]);
                    ou[j] = ub;
                    const float ur = __bfloat162float(ub);
                    s_sum += ur; s_sq += ur * ur;
                }
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const int off = lane * 64 + ((v ^ sw) * 16);
                    *reinterpret_cast<uint4*>(sbuf + off) = reinterpret_cast<const uint4*>(oz)[v];
                    *reinterpret_cast<uint4*>(sbuf + 2048 + off) = reinterpret_cast<const uint4*>(ou)[v];
                }
            } else if (EPI == EPI_POOL_FWD) {
                gelu_fast_n(acc);
                const float cs = warp_transpose_sum(acc, lane);
                colsum[q * kBN + c0 + lane] = cs;
            } else if (EPI == EPI_POOL_DZ) {
                const float dl = __ldg(ep.dpool + (size_t)b * ep.ldn + n0 + c0 + lane);
                __align__(16) bf16 od[32];
                gelu_grad_fast_n(acc);
#pragma unroll
                for (int j = 0; j < 32; ++j) od[j] = __float2bfloat16_rn(__shfl_sync(0xffffffffu, dl, j) * acc[j]);
#pragma unroll
                for (int v = 0; v < 4; ++v)
                    *reinterpret_cast<uint4*>(sbuf + lane * 64 + ((v ^ sw) * 16)) = reinterpret_cast<const uint4*>(od)[v];
            } else if (EPI == EPI_STORE_F32_32) {
                if (ch == 0 && n0 == 0) {
                    float* o = ep.f32out + (size_t)row * 32;
#pragma unroll
                    for (int v = 0; v < 8; ++v)
                        reinterpret_cast<float4*>(o)[v] = make_float4(acc[4 * v], acc[4 * v + 1], acc[4 * v + 2], acc[4 * v + 3]);
                }
            } else if (EPI == EPI_STORE) {
                __align__(16) bf16 oo[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) oo[j] = __float2bfloat16_rn(acc[j]);
#pragma unroll
                for (int v = 0; v < 4; ++v)
                    *reinterpret_cast<uint4*>(sbuf + lane * 64 + ((v ^ sw) * 16)) = reinterpret_cast<const uint4*>(oo)[v];
            } else {
                uint4 bv[4];
#pragma unroll
                for (int v = 0; v < 4; ++v) bv[v] = *reinterpret_cast<const uint4*>(ep.base + grow + c0 + v * 8);
                const bf16* bb = reinterpret_cast<const bf16*>(bv);
                __align__(16) bf16 oo[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) oo[j] = __float2bfloat16_rn(__bfloat162float(bb[j]) + acc[j]);
#pragma unroll
                for (int v = 0; v < 4; ++v)
                    *reinterpret_cast<uint4*>(sbuf + lane * 64 + ((v ^ sw) * 16)) = reinterpret_cast<const uint4*>(oo)[v];
            }
            if (kTmaOut) {
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&mapS0, sbuf, n0 + c0, m0 + q * 32);
                    if (EPI == EPI_LINEAR_FWD) tma_store_2d(&mapS1, sbuf + 2048, n0 + c0, m0 + q * 32);
                    tma_store_commit();
                }
            }
        }
        if (kTmaOut && lane == 0) tma_store_wait<0>();
        if (EPI == EPI_LINEAR_FWD) {
            const double ds = warp_sum_d((double)s_sum), dq = warp_sum_d((double)s_sq);
            if (lane == 0) { atomicAdd(&ep.stats[2 * b], ds); atomicAdd(&ep.stats[2 * b + 1], dq); }
        }
        if (EPI == EPI_POOL_FWD) {
            asm volatile("bar.sync 1, 128;" ::: "memory");           // the four epilogue warps only
            const int t = threadIdx.x - 64;                          // 0..127
            const float tot = colsum[t] + colsum[kBN + t] + colsum[2 * kBN + t] + colsum[3 * kBN + t];
            atomicAdd(&ep.pool[(size_t)b * ep.ldn + n0 + t], tot);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kBN>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------------------
// weight-gradient GEMM: contraction over sequence positions, MN-major operands
// ---------------------------------------------------------------------------------------------------------
struct WgradParams {
    float* out[2]; long stride_m, stride_n, stride_t;  // out[src][m*stride_m + n*stride_n + t*stride_t] += D[m, n]
    int L, dil[2], half, ntaps;                         // blockIdx.y = src * ntaps + t ; shift = (t - half) * dil[src]
    int chunks_total, chunks_per_split;                 // 64-row chunks over all samples
    float* ws;                                          // split-K partial tiles [split][y][tile][128][128], or NULL = atomics
    int head_dv, ntot;                                  // > 0: out is a stack of per-head [ntot, head_dv] matrices (m = head*dv + j)
};
__device__ __forceinline__ size_t wgrad_out_index(const WgradParams& wp, int gm, int gn, int t) {
    if (wp.head_dv > 0)
        return (size_t)(gm / wp.head_dv) * ((size_t)wp.ntot * wp.head_dv) + (size_t)gn * wp.head_dv + (gm % wp.head_dv);
    return (size_t)gm * wp.stride_m + (size_t)gn * wp.stride_n + (size_t)t * wp.stride_t;
}

__global__ void __launch_bounds__(kThreads, 2)
tc_wgrad_kernel(const __grid_constant__ CUtensorMap mapP0, const __grid_constant__ CUtensorMap mapP1,
                const __grid_constant__ CUtensorMap mapQ, WgradParams wp, int tiles_n) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + kStages * kStageBytes);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = (blockIdx.x / tiles_n) * kBM, n0 = (blockIdx.x % tiles_n) * kBN;
    const int src = blockIdx.y / wp.ntaps;
    const int t = blockIdx.y % wp.ntaps;
    const int shift = (t - wp.half) * wp.dil[src];
    const CUtensorMap* mapP = src == 0 ? &mapP0 : &mapP1;
    const int c_beg = blockIdx.z * wp.chunks_per_split;
    const int c_end = min(wp.chunks_total, c_beg + wp.chunks_per_split);
    const int n_iter = c_end - c_beg;
    const int chunks_per_sample = wp.L / 64;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(mapP); tma_prefetch_desc(&mapQ);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<kBN>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    if (n_iter <= 0 && wp.ws != nullptr && warp >= 2) {
        // an empty split still owns a partial tile: write zeros so the reduction can read every slot
        float* dst = wp.ws + ((((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * kBM) * kBN;
        for (int i = threadIdx.x - 64; i < kBM * kBN / 4; i += 128) reinterpret_cast<float4*>(dst)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (n_iter > 0) {
        if (warp == 0) {
            if (lane == 0) {
                for (int i = 0; i < n_iter; ++i) {
                    const int s = i % kStages;
                    mbar_wait(&empty_bar[s], ((i / kStages) & 1) ^ 1);
                    const int chunk = c_beg + i;
                    const int b = chunk / chunks_per_sample;
                    const int l0 = (chunk % chunks_per_sample) * 64;
                    uint8_t* a_dst = smem + s * kStageBytes;
                    uint8_t* b_dst = a_dst + kBM * 64 * 2;
                    mbar_expect_tx(&full_bar[s], kStageBytes);
                    tma_load_3d(a_dst, mapP, &full_bar[s], m0, l0, b);
                    tma_load_3d(a_dst + 8192, mapP, &full_bar[s], m0 + 64, l0, b);
                    tma_load_3d(b_dst, &mapQ, &full_bar[s], n0, l0 + shift, b);
                    tma_load_3d(b_dst + 8192, &mapQ, &full_bar[s], n0 + 64, l0 + shift, b);
                }
            }
        } else if (warp == 1) {
            if (lane == 0) {
                constexpr uint32_t idesc = make_idesc_bf16(kBM, kBN, 1, 1);
                for (int i = 0; i < n_iter; ++i) {
                    const int s = i % kStages;
                    mbar_wait(&full_bar[s], (i / kStages) & 1);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
                    const uint64_t da0 = make_smem_desc_sw128(a_addr, 8192, 1024);
                    const uint64_t db0 = make_smem_desc_sw128(a_addr + kBM * 64 * 2, 8192, 1024);
#pragma unroll
                    for (int k = 0; k < 4; ++k)      // 16 contraction rows per MMA = 2048 B down the tile = +128 in the address field
                        umma_bf16(tmem_base, da0 + (uint64_t)(k * 128), db0 + (uint64_t)(k * 128), idesc, (i == 0 && k == 0) ? 0u : 1u);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(tmem_full_bar);
            }
        } else {
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
            const int q = warp & 3;
            const int m = m0 + q * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
            float* wrow = wp.ws == nullptr ? nullptr
                : wp.ws + ((((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * kBM + (q * 32 + lane)) * kBN;
            float* outp = wp.out[src];
#pragma unroll 1
            for (int ch = 0; ch < kBN / 32; ++ch) {
                float acc[32];
                tmem_ld32(taddr + ch * 32, acc);
                if (wrow != nullptr) {
#pragma unroll
                    for (int v = 0; v < 8; ++v)
                        reinterpret_cast<float4*>(wrow + ch * 32)[v] = make_float4(acc[4 * v], acc[4 * v + 1], acc[4 * v + 2], acc[4 * v + 3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        atomicAdd(outp + wgrad_out_index(wp, m, n0 + ch * 32 + j, t), acc[j]);
                }
            }
            tc_fence_before();
        }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kBN>(tmem_base);
    }
}

// out[src][m, n, t] = sum over splits of the partial tiles (plain stores: every output element is written once).
// block = 16 float4 lanes x 16 split lanes over 64 consecutive partial-tile elements: the loads of one output are
// spread over 16 threads, so a 148-way split costs ~10 dependent loads instead of 148.
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(WgradParams wp, int tiles_m, int tiles_n, int ny, int splits) {
    __shared__ float4 part[16][16];
    pdl_wait();
    const long long total = (long long)ny * tiles_m * tiles_n * kBM * kBN;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const long long base = (long long)blockIdx.x * 64;                 // total is a multiple of 64
    const float4* src = reinterpret_cast<const float4*>(wp.ws + base) + tx;
    const long long stride4 = total / 4;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    int k = ty;
    for (; k + 16 < splits; k += 32) {
        const float4 v0 = src[(long long)k * stride4], v1 = src[(long long)(k + 16) * stride4];
        a.x += v0.x; a.y += v0.y; a.z += v0.z; a.w += v0.w;
        b.x += v1.x; b.y += v1.y; b.z += v1.z; b.w += v1.w;
    }
    if (k < splits) {
        const float4 v0 = src[(long long)k * stride4];
        a.x += v0.x; a.y += v0.y; a.z += v0.z; a.w += v0.w;
    }
    part[ty][tx] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    __syncthreads();
    if (threadIdx.x < 64) {
        const float* pf = reinterpret_cast<const float*>(&part[0][0]);
        float s = 0.f;
#pragma unroll
        for (int y = 0; y < 16; ++y) s += pf[y * 64 + threadIdx.x];
        const long long i = base + threadIdx.x;
        const int n = (int)(i % kBN);
        const int m = (int)((i / kBN) % kBM);
        const int tile = (int)((i / (kBM * kBN)) % (tiles_m * tiles_n));
        const int y = (int)(i / ((long long)kBM * kBN * tiles_m * tiles_n));
        const int srcI = y / wp.ntaps, t = y % wp.ntaps;
        const int gm = (tile / tiles_n) * kBM + m, gn = (tile % tiles_n) * kBN + n;
        wp.out[srcI][wgrad_out_index(wp, gm, gn, t)] = s;
    }
}

int make_mat_map(CUtensorMap* m, const void* base, long rows, int cols, int box_rows) {
    const uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
    const uint64_t strides[1] = {(uint64_t)cols * 2};
    const uint32_t box[2] = {64, (uint32_t)box_rows};
    return pb_make_tmap_bf16(m, base, 2, dims, strides, box);
}
int make_act_map(CUtensorMap* m, const void* base, int B, int L, int C, int box_rows) {
    const uint64_t dims[3] = {(uint64_t)C, (uint64_t)L, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)C * 2, (uint64_t)L * C * 2};
    const uint32_t box[3] = {64, (uint32_t)box_rows, 1};
    return pb_make_tmap_bf16(m, base, 3, dims, strides, box);
}

template <int EPI>
int launch_gemm(const void* A, const void* Bm, const GemmEpi& ep, int M, int N, int K, cudaStream_t st, const char* what) {
    CUtensorMap mA, mB, mS0, mS1;
    int rc;
    if ((rc = make_mat_map(&mA, A, M, K, kBM))) return rc;
    if ((rc = make_mat_map(&mB, Bm, N, K, kBN))) return rc;
    {   // output maps for the TMA-store epilogue: 32 x 32 boxes of the [M, ldn] bf16 outputs, 64-byte swizzle
        const uint64_t dims[2] = {(uint64_t)ep.ldn, (uint64_t)M};
        const uint64_t strides[1] = {(uint64_t)ep.ldn * 2};
        const uint32_t box[2] = {32, 32};
        const void* o0 = ep.out0 ? (const void*)ep.out0 : A;
        const void* o1 = ep.out1 ? (const void*)ep.out1 : o0;
        if ((rc = pb_make_tmap_bf16(&mS0, o0, 2, dims, strides, box, 64))) return rc;
        if ((rc = pb_make_tmap_bf16(&mS1, o1, 2, dims, strides, box, 64))) return rc;
    }
    static bool attr_set = false;
    if (!attr_set) {
        PB_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<EPI>::kSmem));
        attr_set = true;
    }
    PB_CUDA(pb_launch_pdl(tc_gemm_kernel<EPI>, dim3(M / kBM, N / kBN), dim3(kThreads), GemmCfg<EPI>::kSmem, st, mA, mB, mS0, mS1, ep, K));
    pb_count_launch(1);
    return pb_check_launch(what);
}

// P{0,1} [B, L, Mtot], Q [B, L, Ntot]  ->  out{0,1}[m, n, t] (+)= sum_{b,l} P[b,l,m] Q[b,l+shift(t),n]
// With a workspace the split-K partial tiles are stored and reduced by a second kernel (out is overwritten);
// without one they are accumulated with atomics into a zeroed `out`.
int launch_wgrad(const void* P0, const void* P1, const void* Q, float* out0, float* out1, long stride_m, long stride_n,
                 long stride_t, int B, int L, int Mtot, int Ntot, int ntaps, int dil0, int dil1, void* ws, size_t ws_bytes,
                 cudaStream_t st, const char* what, int head_dv = 0) {
    CUtensorMap mP0, mP1, mQ;
    int rc;
    const int nsrc = P1 != nullptr ? 2 : 1;
    if ((rc = make_act_map(&mP0, P0, B, L, Mtot, 64))) return rc;
    if ((rc = make_act_map(&mP1, nsrc == 2 ? P1 : P0, B, L, Mtot, 64))) return rc;
    if ((rc = make_act_map(&mQ, Q, B, L, Ntot, 64))) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        PB_CUDA(cudaFuncSetAttribute(tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        attr_set = true;
    }
    WgradParams wp;
    wp.out[0] = out0; wp.out[1] = out1; wp.stride_m = stride_m; wp.stride_n = stride_n; wp.stride_t = stride_t;
    wp.L = L; wp.dil[0] = dil0; wp.dil[1] = dil1; wp.half = ntaps / 2; wp.ntaps = ntaps;
    wp.head_dv = head_dv; wp.ntot = Ntot;
    wp.chunks_total = B * (L / 64);
    const int tiles_m = Mtot / kBM, tiles_n = Ntot / kBN, ny = nsrc * ntaps;
    const int tiles = tiles_m * tiles_n * ny;
    // one resident wave: about one CTA per SM, at least 8 chunks (512 rows) per CTA
    int splits = 148 / tiles;
    if (splits < 1) splits = 1;
    if (splits > wp.chunks_total / 8) splits = wp.chunks_total / 8 > 0 ? wp.chunks_total / 8 : 1;
    wp.chunks_per_split = (wp.chunks_total + splits - 1) / splits;
    splits = (wp.chunks_total + wp.chunks_per_split - 1) / wp.chunks_per_split;
    const size_t need = (size_t)splits * tiles * kBM * kBN * sizeof(float);
    wp.ws = (ws != nullptr && ws_bytes >= need) ? (float*)ws : nullptr;
    if (wp.ws == nullptr) {           // atomics path needs zeroed outputs
        PB_CUDA(cudaMemsetAsync(out0, 0, (size_t)Mtot * stride_m * sizeof(float), st));
        if (nsrc == 2) PB_CUDA(cudaMemsetAsync(out1, 0, (size_t)Mtot * stride_m * sizeof(float), st));
    }
    PB_CUDA(pb_launch_pdl(tc_wgrad_kernel, dim3(tiles_m * tiles_n, ny, splits), dim3(kThreads), kSmemBytes, st, mP0, mP1, mQ, wp, tiles_n));
    pb_count_launch(1);
    if (wp.ws != nullptr) {
        const long long total = (long long)tiles * kBM * kBN;
        PB_CUDA(pb_launch_pdl(wgrad_reduce_kernel, dim3((unsigned)(total / 64)), dim3(256), 0, st, wp, tiles_m, tiles_n, ny, splits));
        pb_count_launch(1);
    }
    return pb_check_launch(what);
}

}  // namespace

bool pb_tc_linear_ok(int B, int L, int C) { return pb_tc_enabled() && C % 128 == 0 && L % 128 == 0 && B >= 1; }
bool pb_tc_attn_ok(int B, int L, int C, int G) { return pb_tc_enabled() && C % 128 == 0 && G % 128 == 0 && L % 128 == 0 && B >= 1; }

int pb_tc_conv_wgrad(const void* dzn, const void* dzw, const void* x, float* dWn, float* dWw, int B, int L, int C, int ksize,
                     int dil, void* ws, size_t ws_bytes, cudaStream_t st) {
    // dW[co, ci, t] = sum_{b,l} dz[b,l,co] x[b,l+(t-half)*d,ci]; both convolutions (2 x ksize taps) in one launch
    if (pb_tc_mode() != 4 && pb_tc_conv_wgrad3_ok(B, L, C, ksize, dil, ws ? ws_bytes : 0))      // three taps per CTA off one halo tile
        return pb_tc_conv_wgrad3(dzn, dzw, x, dWn, dWw, B, L, ksize, dil, ws, st);
    return launch_wgrad(dzn, dzw, x, dWn, dWw, (long)C * ksize, ksize, 1, B, L, C, C, ksize, 1, dil, ws, ws_bytes, st, "tc_conv_wgrad");
}

int pb_tc_local_linear_fwd(const void* h, const void* W, const float* bias, void* z, void* u2, double* stats2, int B, int L,
                           int C, cudaStream_t st) {
    GemmEpi ep{};
    ep.ldn = C; ep.L = L; ep.bias = bias; ep.base = (const bf16*)h; ep.out0 = (bf16*)z; ep.out1 = (bf16*)u2; ep.stats = stats2;
    ep.base_in_smem = (C == 128) ? 1 : 0;
    return launch_gemm<EPI_LINEAR_FWD>(h, W, ep, B * L, C, C, st, "tc_local_linear_fwd");
}

int pb_tc_local_linear_bwd(const void* du2, const void* dz, const void* h, const void* Wt, void* dh, float* dW, int B, int L,
                           int C, void* ws, size_t ws_bytes, cudaStream_t st) {
    // dh = du2 + dz W   (B operand = W^T [C_in, C_out], K-major)   ;   dW[n, c] = sum_rows dz[row, n] h[row, c]
    GemmEpi ep{};
    ep.ldn = C; ep.L = L; ep.base = (const bf16*)du2; ep.out0 = (bf16*)dh;
    int rc = dh ? launch_gemm<EPI_ADD_STORE>(dz, Wt, ep, B * L, C, C, st, "tc_local_linear_bwd(dh)") : 0;      // dh == NULL: dW only
    if (rc || dW == nullptr) return rc;                                                                        // dW == NULL: dh only
    return launch_wgrad(dz, nullptr, h, dW, nullptr, C, 1, 0, B, L, C, C, 1, 1, 1, ws, ws_bytes, st, "tc_local_linear_bwd(dW)");
}

int pb_tc_attn_pool_fwd(const void* h, const void* wvT, float* pool, int B, int L, int C, int G, cudaStream_t st) {
    GemmEpi ep{};
    ep.ldn = G; ep.L = L; ep.pool = pool;
    if (pb_tc_pool_ok(B, L, C, G)) return pb_tc_pool_fwd(h, wvT, pool, B, L, G, st);      // persistent kernel, weights resident
    return launch_gemm<EPI_POOL_FWD>(h, wvT, ep, B * L, G, C, st, "tc_attn_pool_fwd");
}

int pb_tc_attn_pool_bwd(const void* h, const void* wvT, const void* wv, const float* dpool, const void* dh_in, void* dzv,
                        void* dh_out, float* dWvT, int head_dv, int B, int L, int C, int G, void* ws, size_t ws_bytes, cudaStream_t st) {
    GemmEpi e1{};
    e1.ldn = G; e1.L = L; e1.dpool = dpool; e1.out0 = (bf16*)dzv;
    int rc = 0;
    if (dh_out != nullptr) {        // dh_out == NULL: weight-gradient half only, dzv is then an INPUT
        rc = pb_tc_pool_ok(B, L, C, G) ? pb_tc_pool_dz(h, wvT, dpool, dzv, B, L, G, st)
                                       : launch_gemm<EPI_POOL_DZ>(h, wvT, e1, B * L, G, C, st, "tc_attn_pool_bwd(dzv)");
        if (rc) return rc;
        GemmEpi e2{};
        e2.ldn = C; e2.L = L; e2.base = (const bf16*)dh_in; e2.out0 = (bf16*)dh_out;
        rc = launch_gemm<EPI_ADD_STORE>(dzv, wv, e2, B * L, C, G, st, "tc_attn_pool_bwd(dh)");
        if (rc) return rc;
    }
    if (dWvT) rc = launch_wgrad(dzv, nullptr, h, dWvT, nullptr, C, 1, 0, B, L, G, C, 1, 1, 1, ws, ws_bytes, st, "tc_attn_pool_bwd(dWv)", head_dv);
    return rc;
}

// ---- token head on tensor cores (modules.py:277-284): logits = h Wo^T with Wo zero-padded to 128 rows ----------
bool pb_tc_token_ok(int B, int L, int C, int V) { return pb_tc_enabled() && C % 128 == 0 && L % 128 == 0 && V <= 32 && B >= 1; }

int pb_tc_token_logits(const void* h, const void* wo_pad, float* logits32, int B, int L, int C, cudaStream_t st) {
    GemmEpi ep{};
    ep.ldn = 128; ep.L = L; ep.f32out = logits32;
    return launch_gemm<EPI_STORE_F32_32>(h, wo_pad, ep, B * L, 128, C, st, "tc_token_logits");
}

int pb_tc_token_bwd(const void* dlogit_pad, const void* woT_pad, const void* h, void* dh, float* dWo_pad, int B, int L, int C,
                    void* ws, size_t ws_bytes, cudaStream_t st) {
    // dh = dlogit Wo (K = 128 padded vocabulary)  ;  dWo_pad[v, c] = sum_rows dlogit[row, v] h[row, c]
    // either half may be skipped (NULL output): callers put the weight gradient on a second stream
    if (dh != nullptr) {
        GemmEpi ep{};
        ep.ldn = C; ep.L = L; ep.out0 = (bf16*)dh;
        int rc = launch_gemm<EPI_STORE>(dlogit_pad, woT_pad, ep, B * L, C, 128, st, "tc_token_bwd(dh)");
        if (rc) return rc;
    }
    if (dWo_pad == nullptr) return PB_OK;
    return launch_wgrad(dlogit_pad, nullptr, h, dWo_pad, nullptr, C, 1, 0, B, L, 128, C, 1, 1, 1, ws, ws_bytes, st, "tc_token_bwd(dWo)");
}
