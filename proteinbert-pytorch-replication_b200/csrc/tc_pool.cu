// Attention pooling GEMMs as ONE persistent CTA per SM (modules.py:21-92 under the reference's softmax axis):
//   POOL_FWD   pool[b, n] += sum over the tile's rows of gelu(h Wv)          (forward)
//   POOL_DZ    dzv = dpool[b, n] * gelu'(h Wv)                               (backward, first GEMM)
// K = C = 128 only: eight MMAs per 128 x 128 tile against 16 k GELU evaluations, so these kernels are bound by the FP32
// pipe in their epilogues (profiles/r01_tc_gemm_ncu.md).  With one tile per CTA the epilogue warps idle while the next
// CTA sets up (barriers, TMEM allocation, TMA round trip, MMAs).  Here the whole weight matrix Wv^T [G x 128] stays in
// shared memory (128 KB), the CTA walks (row tile, column tile) units, the MMA warp runs up to four units ahead into
// four TMEM accumulators (all 512 columns) and four groups of four epilogue warps drain one accumulator each, so the
// FP32 pipe never waits for a tile and every scheduler has four epilogue warps to pick from.  ncu (round 2,
// profiles/r02_conv_pool_ncu.md): issue slots 62-68 % busy, FMA pipe 39-42 %, ALU pipe 26-39 %: instruction count per
// element is what is left to cut.
#include "tc_common.cuh"
#include "tc_api.cuh"

namespace {

constexpr int kTile = 128 * 64 * 2;                  // one 128-row x 64-k bf16 tile: 16 KB
constexpr int kAStages = 3;                          // A ring: stage = one 128 x 128 activation tile (two k-chunks)
constexpr int kMaxTilesN = 8;                        // column tiles of the weights (G <= 1024); a CTA keeps ONE of them resident (32 KB)
constexpr int kGroups = 4;                           // epilogue groups of four warps, one per TMEM accumulator
constexpr int kStageArea = kGroups * 4 * 2 * 2048;   // POOL_DZ: two 2 KB staging tiles per epilogue warp
constexpr int kSmemBytes = 2 * kTile + kAStages * 2 * kTile + kStageArea + 1024 + 1024;   // barriers + TMEM slot in the last KB
constexpr int kThreads = 64 + kGroups * 128;         // warp 0 TMA, warp 1 MMA + TMEM, then the epilogue groups

enum { POOL_FWD = 0, POOL_DZ = 1 };

struct PoolParams {
    int L, G;                       // rows per sample, total columns
    int tiles_n, row_tiles;         // units = row tiles x column tiles; CTA c owns column tile c % tiles_n and walks row tiles
                                    // c / tiles_n, c / tiles_n + gridDim.x / tiles_n, ... (gridDim.x is a multiple of tiles_n)
    float* pool;                    // POOL_FWD [B, G] (accumulated)
    const float* dpool;             // POOL_DZ  [B, G]
};

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
tc_pool_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
               const __grid_constant__ CUtensorMap mapS, PoolParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* b_res = smem;                                           // [2][16 KB]: this CTA's column tile of the weights
    uint8_t* a_ring = smem + 2 * kTile;                              // [kAStages][2][16 KB]
    uint8_t* stage_area = a_ring + kAStages * 2 * kTile;
    uint64_t* b_full = (uint64_t*)(stage_area + kStageArea);
    uint64_t* a_full = b_full + 1;
    uint64_t* a_empty = a_full + kAStages;
    uint64_t* acc_full = a_empty + kAStages;                         // [4]
    uint64_t* acc_empty = acc_full + 4;                              // [4]
    uint32_t* tmem_slot = (uint32_t*)(acc_empty + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // this CTA's units: column tile n_cta, row tiles m0, m0 + m_step, ...  (One resident column tile instead of all of them:
    // 32 KB instead of 128 KB to wait for before the first MMA, and room for a third activation stage and double-buffered
    // output staging.  The activation tile is fetched per unit either way.)
    const int n_cta = (int)blockIdx.x % p.tiles_n, m0 = (int)blockIdx.x / p.tiles_n, m_step = (int)gridDim.x / p.tiles_n;
    const int n_mine = p.row_tiles > m0 ? (p.row_tiles - m0 + m_step - 1) / m_step : 0;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA); tma_prefetch_desc(&mapB);
        mbar_init(b_full, 1);
        for (int s = 0; s < kAStages; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
        for (int a = 0; a < 4; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 4); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    if (warp == 0) {
        if (lane == 0 && n_mine > 0) {
            mbar_expect_tx(b_full, 2 * kTile);
            for (int kc = 0; kc < 2; ++kc) tma_load_2d(b_res + kc * kTile, &mapB, b_full, kc * 64, n_cta * 128);
            for (int i = 0; i < n_mine; ++i) {
                const int s = i % kAStages;
                const int m = m0 + i * m_step;
                mbar_wait(&a_empty[s], ((i / kAStages) & 1) ^ 1);
                mbar_expect_tx(&a_full[s], 2 * kTile);
                tma_load_2d(a_ring + (s * 2) * kTile, &mapA, &a_full[s], 0, m * 128);
                tma_load_2d(a_ring + (s * 2 + 1) * kTile, &mapA, &a_full[s], 64, m * 128);
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && n_mine > 0) {
            constexpr uint32_t idesc = make_idesc_bf16(128, 128, 0, 0);
            mbar_wait(b_full, 0);
            for (int i = 0; i < n_mine; ++i) {
                const int s = i % kAStages, a = i & 3;
                mbar_wait(&acc_empty[a], ((i >> 2) & 1) ^ 1);        // the epilogue has drained this accumulator
                mbar_wait(&a_full[s], (i / kAStages) & 1);
                tc_fence_after();
#pragma unroll
                for (int kc = 0; kc < 2; ++kc) {
                    const uint64_t da = make_smem_desc_sw128(smem_u32(a_ring + (s * 2 + kc) * kTile), 16, 1024);
                    const uint64_t db = make_smem_desc_sw128(smem_u32(b_res + kc * kTile), 16, 1024);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        // POOL_FWD: weights as the A operand (M = 128 columns of Wv), positions as N -- accumulator lane = column
                        // of Wv; POOL_DZ: positions as M (its output rows are positions)
                        umma_bf16(tmem_base + a * 128, (EPI == POOL_FWD ? db : da) + (uint64_t)(k * 2),
                                  (EPI == POOL_FWD ? da : db) + (uint64_t)(k * 2), idesc, (kc == 0 && k == 0) ? 0u : 1u);
                }
                umma_commit(&a_empty[s]);
                umma_commit(&acc_full[a]);
            }
        }
    } else {
        // ===== epilogue: groups of four warps; group g drains units g, g + kGroups, ... (always accumulator g).
        //       POOL_FWD accumulators are TRANSPOSED (TMEM lane = column of Wv, TMEM column = position of the row tile): the sum
        //       over positions is then a sum inside each thread.  (With positions as lanes the column sums took a 32 x 32
        //       shuffle transpose per chunk: 7.6 of the 24.5 instructions per element; 16.8 -> 14.4 us per launch.)
        //       POOL_DZ keeps positions as lanes: its output rows are positions, so 16-byte row stores beat the 2-byte
        //       transposing stores (17.3 against 17.8 us), even with one shuffle per element to broadcast dpool.
        //       Four groups = four warps per scheduler: these epilogues are bound by instruction issue =====
        const int grp = (warp - 2) >> 2, q = warp & 3;
        uint8_t* stage0 = stage_area + (warp - 2) * 2 * 2048;                        // POOL_DZ: two staging tiles per warp
        const int sw = (lane >> 1) & 3;
        int n_staged = 0;
#pragma unroll 1
        for (int i = grp; i < n_mine; i += kGroups) {
            const int a = i & 3;
            const int m = m0 + i * m_step, n = n_cta;
            const int b = (m * 128) / p.L;                           // tiles never straddle samples (L % 128 == 0)
            mbar_wait(&acc_full[a], (i >> 2) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * 128;
            float colsum = 0.f;                                      // POOL_FWD: column n*128 + q*32 + lane over the tile's positions
#pragma unroll 1
            for (int ch = 0; ch < 4; ++ch) {
                float acc[32];
                tmem_ld32(taddr + ch * 32, acc);
                const int c0 = ch * 32;
                if (ch == 3) {                                       // last read of this accumulator: hand it back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[a]);
                }
                if (EPI == POOL_FWD) {
                    float s0 = 0.f, s1 = 0.f;
                    gelu_fast_n(acc);
#pragma unroll
                    for (int j = 0; j < 32; j += 2) { s0 += acc[j]; s1 += acc[j + 1]; }
                    colsum += s0 + s1;
                } else {
                    const float dl = __ldg(p.dpool + (size_t)b * p.G + n * 128 + c0 + lane);
                    __align__(16) bf16 od[32];
                    gelu_grad_fast_n(acc);
#pragma unroll
                    for (int j = 0; j < 32; ++j) od[j] = __float2bfloat16_rn(__shfl_sync(0xffffffffu, dl, j) * acc[j]);
                    uint8_t* stage = stage0 + (n_staged & 1) * 2048;
                    if (n_staged >= 2) {                             // the buffer's previous store (two stores ago) must have read it
                        if (lane == 0) tma_store_wait_read<1>();
                        __syncwarp();
                    }
                    ++n_staged;
#pragma unroll
                    for (int v = 0; v < 4; ++v)
                        *reinterpret_cast<uint4*>(stage + lane * 64 + ((v ^ sw) * 16)) = reinterpret_cast<const uint4*>(od)[v];
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) { tma_store_2d(&mapS, stage, n * 128 + c0, m * 128 + q * 32); tma_store_commit(); }
                }
            }
            if (EPI == POOL_FWD) atomicAdd(&p.pool[(size_t)b * p.G + n * 128 + q * 32 + lane], colsum);
        }
        if (EPI == POOL_DZ && lane == 0) tma_store_wait<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

int mat_map(CUtensorMap* m, const void* base, long rows, int cols) {
    const uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
    const uint64_t strides[1] = {(uint64_t)cols * 2};
    const uint32_t box[2] = {64, 128};
    return pb_make_tmap_bf16(m, base, 2, dims, strides, box);
}

template <int EPI>
int launch_pool(const void* h, const void* wvT, void* dzv, PoolParams p, int M, cudaStream_t st, const char* what) {
    CUtensorMap mA, mB, mS;
    int rc;
    if ((rc = mat_map(&mA, h, M, 128))) return rc;
    if ((rc = mat_map(&mB, wvT, p.G, 128))) return rc;
    {
        const uint64_t dims[2] = {(uint64_t)p.G, (uint64_t)M};
        const uint64_t strides[1] = {(uint64_t)p.G * 2};
        const uint32_t box[2] = {32, 32};
        if ((rc = pb_make_tmap_bf16(&mS, dzv ? dzv : h, 2, dims, strides, box, 64))) return rc;
    }
    static bool attr_set = false;
    if (!attr_set) {
        PB_CUDA(cudaFuncSetAttribute(tc_pool_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        attr_set = true;
    }
    int grid = (148 / p.tiles_n) * p.tiles_n;                        // a multiple of tiles_n: every CTA owns one column tile
    if (grid > p.row_tiles * p.tiles_n) grid = p.row_tiles * p.tiles_n;
    PB_CUDA(pb_launch_pdl(tc_pool_kernel<EPI>, dim3(grid), dim3(kThreads), kSmemBytes, st, mA, mB, mS, p));
    pb_count_launch(1);
    return pb_check_launch(what);
}

}  // namespace

bool pb_tc_pool_ok(int B, int L, int C, int G) {
    return pb_tc_enabled() && pb_tc_mode() != 4 && C == 128 && G % 128 == 0 && G / 128 <= kMaxTilesN && L % 128 == 0 && B >= 1;
}

int pb_tc_pool_fwd(const void* h, const void* wvT, float* pool, int B, int L, int G, cudaStream_t st) {
    PoolParams p{};
    p.L = L; p.G = G; p.tiles_n = G / 128; p.row_tiles = B * L / 128; p.pool = pool;
    return launch_pool<POOL_FWD>(h, wvT, nullptr, p, B * L, st, "tc_pool_fwd");
}

int pb_tc_pool_dz(const void* h, const void* wvT, const float* dpool, void* dzv, int B, int L, int G, cudaStream_t st) {
    PoolParams p{};
    p.L = L; p.G = G; p.tiles_n = G / 128; p.row_tiles = B * L / 128; p.dpool = dpool;
    return launch_pool<POOL_DZ>(h, wvT, dzv, p, B * L, st, "tc_pool_dz");
}
