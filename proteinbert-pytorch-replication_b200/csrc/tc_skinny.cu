// The annotation (GO) layers on tcgen05: Linear(A -> G) at the input and Linear(G -> A) at the output of the model
// (modules.py:248-251, 290-297) and their backward.  With M = batch rows <= 128 these are "skinny" GEMMs: 0.6 GFLOP
// against an 18 MB fp32 weight matrix each, i.e. bound by streaming the weights once.  The fp32 SIMT engine is
// FMA-bound on them (about 30 us each at batch 64); here the MMA cost vanishes and the kernels run at the speed the
// producer warps can pull the weights in.
//
// Operands stay the caller's fp32 tensors, with any leading dimension (A = 8943 is odd, so TMA cannot describe
// them): eight producer warps read them with coalesced loads, round to bf16 and write the 128-byte-swizzled tile
// layouts the MMA descriptors expect (generic-proxy stores, then fence.proxy.async + mbarrier arrive); warp 8
// issues tcgen05.mma into a 128 x 128 fp32 TMEM accumulator.  Three operand arrangements cover the five products:
//   forward   Y  = X W^T          A K-major (X rows)        B K-major (W rows)          split over K
//   backward  dX = dZ W           A K-major (dZ rows)       B MN-major (W rows are K)   split over N
//   backward  dW = dZ^T X         A MN-major (dZ rows = K)  B MN-major (X rows = K)     contraction = batch, no split
// Accumulators leave through a per-warp shared-memory transpose so every global store is a coalesced 128-byte row
// segment, whatever the alignment of the output rows.
#include "tc_common.cuh"
#include "tc_api.cuh"
#include "../../include/pb_b200.h"

namespace {

constexpr int kTileBytes = 128 * 64 * 2;              // one operand tile: 128 x 64 bf16
constexpr int kStages = 3;
constexpr int kStageBytes = 2 * kTileBytes;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
constexpr int kProducers = 256;                     // 8 producer warps; warp 8 issues the MMAs
constexpr int kThreads = 288;

struct SkinnyParams {
    const float* A; long lda; int a_valid;            // a_valid: extent of A's M index (rows of the output)
    const float* B; long ldb; int b_valid;            // b_valid: extent of B's N index (columns of the output)
    int K;                                            // contraction extent
    int chunks_per_split;                             // 64-wide contraction chunks per blockIdx.z
    float* out; long ldo; long split_stride;          // out[z * split_stride + m * ldo + n]
};

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// One 128 x 64 operand tile, fp32 global -> bf16 shared, 128-byte rows with the 16-byte chunks XOR-swizzled by
// (row & 7).  K-major: shared row r = M/N index mn0 + r, its 64 elements = contraction indices k0.. ;
// MN-major: two blocks of 64 rows, row kk of block nb = contraction index k0 + kk, elements = M/N indices
// mn0 + 64 nb .. ; either way a shared row is 64 consecutive floats in global memory and a warp copies it with
// two coalesced loads.  Anything outside [0, mn_end) x [0, k_end) is written as zero.
template <bool MN>
__device__ __forceinline__ void tile_row(const float* __restrict__ p, long ld, int mn0, int mn_end, int k0, int k_end, int row, int lane,
                                         float& v0, float& v1) {
    const float* base;
    bool ok;
    int lim;
    if (MN) {
        const int nb = row >> 6, gk = k0 + (row & 63);
        ok = gk < k_end;
        base = p + (size_t)gk * ld + mn0 + nb * 64;
        lim = mn_end - (mn0 + nb * 64);
    } else {
        const int gr = mn0 + row;
        ok = gr < mn_end;
        base = p + (size_t)gr * ld + k0;
        lim = k_end - k0;
    }
    const int e = 2 * lane;
    v0 = (ok && e < lim) ? __ldg(base + e) : 0.f;
    v1 = (ok && e + 1 < lim) ? __ldg(base + e + 1) : 0.f;
}
__device__ __forceinline__ void tile_store(uint8_t* dst, int row, int lane, float v0, float v1) {
    *reinterpret_cast<uint32_t*>(dst + row * 128 + ((((lane >> 2) ^ (row & 7))) << 4) + (lane & 3) * 4) = pack_bf16x2(v0, v1);
}
// Each of the 8 producer warps owns 16 rows of the A tile and 16 rows of the B tile; all 64 loads of a chunk are
// issued before the first store, so a CTA keeps 64 KB of weight/activation bytes in flight.
template <bool A_MN, bool B_MN>
__device__ __forceinline__ void load_chunk(uint8_t* a_dst, uint8_t* b_dst, const float* __restrict__ A, long lda, int m0, int a_end,
                                           const float* __restrict__ B, long ldb, int n0, int b_end, int k0, int k_end, int warp,
                                           int lane) {
    float va[16][2], vb[16][2];
#pragma unroll
    for (int i = 0; i < 16; ++i) tile_row<B_MN>(B, ldb, n0, b_end, k0, k_end, warp * 16 + i, lane, vb[i][0], vb[i][1]);
#pragma unroll
    for (int i = 0; i < 16; ++i) tile_row<A_MN>(A, lda, m0, a_end, k0, k_end, warp * 16 + i, lane, va[i][0], va[i][1]);
#pragma unroll
    for (int i = 0; i < 16; ++i) tile_store(b_dst, warp * 16 + i, lane, vb[i][0], vb[i][1]);
#pragma unroll
    for (int i = 0; i < 16; ++i) tile_store(a_dst, warp * 16 + i, lane, va[i][0], va[i][1]);
}

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kThreads, 2) tc_skinny_kernel(SkinnyParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + kStages * kStageBytes);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * 128, n0 = blockIdx.y * 128;
    const int chunks_total = (p.K + 63) / 64;
    const int c_beg = blockIdx.z * p.chunks_per_split;
    const int c_end = min(chunks_total, c_beg + p.chunks_per_split);
    const int n_iter = c_end - c_beg;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], kProducers); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 8) tmem_alloc<128>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    if (warp < 8) {
        for (int i = 0; i < n_iter; ++i) {
            const int s = i % kStages;
            mbar_wait(&empty_bar[s], ((i / kStages) & 1) ^ 1);
            const int k0 = (c_beg + i) * 64;
            uint8_t* a_dst = smem + s * kStageBytes;
            load_chunk<A_MN, B_MN>(a_dst, a_dst + kTileBytes, p.A, p.lda, m0, p.a_valid, p.B, p.ldb, n0, p.b_valid, k0, p.K, warp, lane);
            fence_proxy_async();                      // make the generic-proxy stores visible to the tensor core
            mbar_arrive(&full_bar[s]);
        }
        // ---- epilogue: TMEM lane = output row; transpose 32 x 32 blocks through shared memory -----------------
        float* tb = reinterpret_cast<float*>(smem) + warp * (32 * 33);       // the ring is idle once the accumulator is complete
        if (n_iter > 0) {
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
        }
        const int q = warp & 3;                                              // TMEM lane quadrant this warp may read
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        float* outz = p.out + (size_t)blockIdx.z * p.split_stride;
#pragma unroll 1
        for (int ch = (warp >> 2) * 2; ch < (warp >> 2) * 2 + 2; ++ch) {     // warps 0-3: columns 0-63, warps 4-7: 64-127
            float acc[32];
            if (n_iter > 0) {
                tmem_ld32(taddr + ch * 32, acc);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) acc[j] = 0.f;       // an empty split still owns (zero) partial sums
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) tb[lane * 33 + j] = acc[j];
            __syncwarp();
            const int n = n0 + ch * 32 + lane;
            if (n < p.b_valid) {
#pragma unroll 8
                for (int r = 0; r < 32; ++r) {
                    const int m = m0 + q * 32 + r;
                    if (m < p.a_valid) outz[(size_t)m * p.ldo + n] = tb[r * 33 + lane];
                }
            }
            __syncwarp();
        }
        tc_fence_before();
    } else if (lane == 0) {
        constexpr uint32_t idesc = make_idesc_bf16(128, 128, A_MN ? 1 : 0, B_MN ? 1 : 0);
        for (int i = 0; i < n_iter; ++i) {
            const int s = i % kStages;
            mbar_wait(&full_bar[s], (i / kStages) & 1);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
            const uint64_t da0 = A_MN ? make_smem_desc_sw128(a_addr, 8192, 1024) : make_smem_desc_sw128(a_addr, 16, 1024);
            const uint64_t db0 = B_MN ? make_smem_desc_sw128(a_addr + kTileBytes, 8192, 1024) : make_smem_desc_sw128(a_addr + kTileBytes, 16, 1024);
#pragma unroll
            for (int k = 0; k < 4; ++k)        // 16 contraction indices: +32 B along a K-major row, +16 rows (2048 B) down an MN-major tile
                umma_bf16(tmem_base, da0 + (uint64_t)(k * (A_MN ? 128 : 2)), db0 + (uint64_t)(k * (B_MN ? 128 : 2)), idesc,
                          (i == 0 && k == 0) ? 0u : 1u);
            umma_commit(&empty_bar[s]);
        }
        if (n_iter > 0) umma_commit(tmem_full_bar);
    }
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc<128>(tmem_base);
    }
}

// out[i] (+)= act(sum over splits + bias): block = `per` outputs x (256 / per) split lanes
__global__ void __launch_bounds__(256) skinny_reduce_kernel(const float* __restrict__ ws, const float* __restrict__ bias, float* Z,
                                                            float* Y, long long total, int N, int nsplit, int act, int accumulate,
                                                            int per) {
    __shared__ float part[256];
    pdl_wait();
    const int lanes = 256 / per;
    const int tx = threadIdx.x % per, ty = threadIdx.x / per;
    const long long i = (long long)blockIdx.x * per + tx;
    float s = 0.f;
    if (i < total) {
        int k = ty;
        for (; k + 3 * lanes < nsplit; k += 4 * lanes) {
            const float a = ws[(size_t)k * total + i], b = ws[(size_t)(k + lanes) * total + i];
            const float c = ws[(size_t)(k + 2 * lanes) * total + i], d = ws[(size_t)(k + 3 * lanes) * total + i];
            s += (a + b) + (c + d);
        }
        for (; k < nsplit; k += lanes) s += ws[(size_t)k * total + i];
    }
    part[threadIdx.x] = s;
    __syncthreads();
    if (ty == 0 && i < total) {
        for (int y = 1; y < lanes; ++y) s += part[y * per + tx];
        const float z = s + (bias ? bias[i % N] : 0.f);
        if (Z) Z[i] = z;
        const float y = act_fwd(act, z);
        Y[i] = accumulate ? Y[i] + y : y;
    }
}

// dZ = dY * act'(ZY); db[n] = sum_m dZ[m, n].  block = 32 columns x 8 row lanes
__global__ void skinny_dz_kernel(const float* __restrict__ dY, const float* __restrict__ ZY, float* __restrict__ dZ,
                                 float* __restrict__ db, int M, int N, int act) {
    __shared__ float part[8][33];
    pdl_wait();
    const int n = blockIdx.x * 32 + threadIdx.x;
    float s = 0.f;
    if (n < N) {
        for (int m = threadIdx.y; m < M; m += 8) {
            const size_t i = (size_t)m * N + n;
            const float g = dY[i] * (act == PB_ACT_NONE ? 1.f : act_bwd(act, ZY[i]));
            dZ[i] = g;
            s += g;
        }
    }
    part[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && n < N && db) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += part[k][threadIdx.x];
        db[n] = t;
    }
}

template <bool A_MN, bool B_MN>
int launch_skinny(const SkinnyParams& p, int tiles_m, int tiles_n, int splits, cudaStream_t st) {
    static bool attr_set = false;
    if (!attr_set) {
        PB_CUDA(cudaFuncSetAttribute(tc_skinny_kernel<A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        attr_set = true;
    }
    PB_CUDA(pb_launch_pdl(tc_skinny_kernel<A_MN, B_MN>, dim3(tiles_m, tiles_n, splits), dim3(kThreads), kSmemBytes, st, p));
    pb_count_launch(1);
    return PB_OK;
}

// splits over the contraction so that about two CTAs per SM are in flight
void pick_splits(int tiles, int K, int* splits, int* chunks_per_split) {
    const int chunks = (K + 63) / 64;
    int s = 296 / (tiles > 0 ? tiles : 1);            // two CTAs per SM
    if (s < 1) s = 1;
    if (s > chunks) s = chunks;
    const int cps = (chunks + s - 1) / s;
    *chunks_per_split = cps;
    *splits = (chunks + cps - 1) / cps;
}

int launch_reduce(const float* ws, const float* bias, float* Z, float* Y, long long total, int N, int nsplit, int act, int accumulate,
                  cudaStream_t st) {
    const int per = nsplit >= 8 ? 32 : 256;
    PB_CUDA(pb_launch_pdl(skinny_reduce_kernel, dim3((unsigned)((total + per - 1) / per)), dim3(256), 0, st, ws, bias, Z, Y, total, N,
                          nsplit, act, accumulate, per));
    pb_count_launch(1);
    return PB_OK;
}

}  // namespace

extern "C" {

int pb_linear_tc_ok(int M, int N, int K) { return pb_tc_enabled() && M >= 1 && M <= 128 && N >= 1 && K >= 1; }

size_t pb_linear_tc_workspace(int M, int N, int K) {
    // forward: splits x M x N ; backward dX: splits x M x K ; splits <= 296 / tiles, tiles >= ceil(extent / 128)
    auto need = [&](int cols, int contr) {
        int s, cps;
        pick_splits((cols + 127) / 128, contr, &s, &cps);
        return (size_t)s * M * cols * sizeof(float);
    };
    const size_t a = need(N, K), b = need(K, N);
    return a > b ? a : b;
}

int pb_linear_fwd_tc(const float* X, const float* W, const float* bias, float* Z, float* Y, int M, int N, int K, int act, void* ws,
                     size_t ws_bytes, void* stream) {
    PB_REQUIRE(pb_linear_tc_ok(M, N, K), PB_ERR_SHAPE, "linear_fwd_tc: needs 1 <= M <= 128 (got M=%d N=%d K=%d) and the tensor-core engine on", M, N, K);
    cudaStream_t st = (cudaStream_t)stream;
    const int tiles_n = (N + 127) / 128;
    int splits, cps;
    pick_splits(tiles_n, K, &splits, &cps);
    PB_REQUIRE(ws != nullptr && ws_bytes >= (size_t)splits * M * N * sizeof(float), PB_ERR_WORKSPACE,
               "linear_fwd_tc: workspace of %zu bytes required (pb_linear_tc_workspace)", (size_t)splits * M * N * sizeof(float));
    SkinnyParams p;
    p.A = X; p.lda = K; p.a_valid = M;
    p.B = W; p.ldb = K; p.b_valid = N;
    p.K = K; p.chunks_per_split = cps;
    p.out = (float*)ws; p.ldo = N; p.split_stride = (long)M * N;
    int rc = launch_skinny<false, false>(p, 1, tiles_n, splits, st);
    if (rc) return rc;
    rc = launch_reduce((const float*)ws, bias, Z, Y, (long long)M * N, N, splits, act, 0, st);
    if (rc) return rc;
    return pb_check_launch("pb_linear_fwd_tc");
}

int pb_linear_bwd_tc(const float* dY, const float* ZY, const float* X, const float* W, float* dZ_buf, float* dX, float* dW, float* db,
                     int M, int N, int K, int act, int accumulate_dx, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(pb_linear_tc_ok(M, N, K), PB_ERR_SHAPE, "linear_bwd_tc: needs 1 <= M <= 128 (got M=%d N=%d K=%d) and the tensor-core engine on", M, N, K);
    PB_REQUIRE(dZ_buf != nullptr, PB_ERR_WORKSPACE, "linear_bwd_tc: dZ_buf is required");
    PB_REQUIRE(dY == nullptr || act == PB_ACT_NONE || ZY != nullptr, PB_ERR_SHAPE, "linear_bwd_tc: ZY required for act=%d", act);
    cudaStream_t st = (cudaStream_t)stream;
    if (dY != nullptr) {       // dY == NULL: dZ_buf already holds dZ from an earlier call (and db is done)
        PB_CUDA(pb_launch_pdl(skinny_dz_kernel, dim3((N + 31) / 32), dim3(32, 8), 0, st, dY, ZY, dZ_buf, db, M, N, act));
        pb_count_launch(1);
    }
    int rc;
    if (dX) {       // dX[m, k] = sum_n dZ[m, n] W[n, k]
        const int tiles_k = (K + 127) / 128;
        int splits, cps;
        pick_splits(tiles_k, N, &splits, &cps);
        PB_REQUIRE(ws != nullptr && ws_bytes >= (size_t)splits * M * K * sizeof(float), PB_ERR_WORKSPACE,
                   "linear_bwd_tc: workspace of %zu bytes required (pb_linear_tc_workspace)", (size_t)splits * M * K * sizeof(float));
        SkinnyParams p;
        p.A = dZ_buf; p.lda = N; p.a_valid = M;
        p.B = W; p.ldb = K; p.b_valid = K;
        p.K = N; p.chunks_per_split = cps;
        p.out = (float*)ws; p.ldo = K; p.split_stride = (long)M * K;
        if ((rc = launch_skinny<false, true>(p, 1, tiles_k, splits, st))) return rc;
        if ((rc = launch_reduce((const float*)ws, nullptr, nullptr, dX, (long long)M * K, K, splits, PB_ACT_NONE, accumulate_dx, st))) return rc;
    }
    if (dW) {       // dW[n, k] = sum_m dZ[m, n] X[m, k]: the contraction is the batch
        SkinnyParams p;
        p.A = dZ_buf; p.lda = N; p.a_valid = N;
        p.B = X; p.ldb = K; p.b_valid = K;
        p.K = M; p.chunks_per_split = (M + 63) / 64;
        p.out = dW; p.ldo = K; p.split_stride = 0;
        if ((rc = launch_skinny<true, true>(p, (N + 127) / 128, (K + 127) / 128, 1, st))) return rc;
    }
    return pb_check_launch("pb_linear_bwd_tc");
}

}  // extern "C"
