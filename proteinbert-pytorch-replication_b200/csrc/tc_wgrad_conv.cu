// Weight gradients of the two convolutions of a block (modules.py:133-147, run in reverse) on tcgen05:
//   dW[src][co, ci, t] = sum_{b,l} dz_src[b, l, co] * x[b, l + (t - half) * dil_src, ci]
//
// The general kernel in tc_gemm.cu gives every tap its own CTA, which streams dz and x from L2 once per tap
// (18 x 16 MB per block at B=64, L=512) and is bound by that traffic.  Here one CTA owns THREE adjacent taps of one
// convolution: per 64-position chunk it fetches the dz tile once and one x halo tile of 64 + 2*dil rows, and feeds
// the three taps to three TMEM accumulators from row-shifted views of the same halo (the contraction runs over
// positions, so a tap shift is a shift of the MN-major operand's start address by dil rows of 128 bytes; the 128-byte
// swizzle is a function of the absolute shared-memory address, so the shifted descriptor keeps base_offset = 0).
// Operand traffic per MMA drops 2.7x; partial tiles go to the workspace and a second kernel sums the splits and
// writes the [co, ci, t] layout with coalesced stores.
//
// grid = (ksize / 3 tap groups, 2 convolutions, splits over position chunks) ~ one CTA per SM (512 TMEM columns).
#include "tc_common.cuh"
#include "tc_api.cuh"

namespace {

constexpr int kTaps = 3;                              // taps (accumulators) per CTA
constexpr int kRegion = 10240;                        // one 64-channel block of the halo: up to 80 rows of 128 B
constexpr int kDzBytes = 2 * 64 * 128;                // 64 positions x 128 channels
constexpr int kStageBytes = kDzBytes + 2 * kRegion;   // 36 KB
constexpr int kStages = 5;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
constexpr int kThreads = 192;
constexpr int kTile = 128 * 128;

struct ConvWgradParams {
    int L, ksize, half, dil[2];
    int chunks_total, chunks_per_split;
    float* ws;                                        // [split][src][tap][128][128] partial tiles
};

__device__ __forceinline__ uint64_t desc_mn_shifted(uint32_t saddr, uint32_t lbo) {
    return make_smem_desc_sw128(saddr, lbo, 1024) & ~((uint64_t)7 << 49);
}

__global__ void __launch_bounds__(kThreads, 1)
tc_conv_wgrad3_kernel(const __grid_constant__ CUtensorMap mapP0, const __grid_constant__ CUtensorMap mapP1,
                      const __grid_constant__ CUtensorMap mapX0, const __grid_constant__ CUtensorMap mapX1, ConvWgradParams wp) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + kStages * kStageBytes);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tg = blockIdx.x, src = blockIdx.y;
    const int d = wp.dil[src];
    const int shift0 = (tg * kTaps - wp.half) * d;            // halo row 0 sits at position l0 + shift0
    const int halo_rows = 64 + (kTaps - 1) * d;
    const CUtensorMap* mapP = src == 0 ? &mapP0 : &mapP1;
    const CUtensorMap* mapX = src == 0 ? &mapX0 : &mapX1;
    const int c_beg = blockIdx.z * wp.chunks_per_split;
    const int c_end = min(wp.chunks_total, c_beg + wp.chunks_per_split);
    const int n_iter = c_end - c_beg;
    const int chunks_per_sample = wp.L / 64;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(mapP); tma_prefetch_desc(mapX);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    float* part = wp.ws + (((size_t)blockIdx.z * gridDim.y + src) * wp.ksize + tg * kTaps) * kTile;
    if (n_iter <= 0) {
        // an empty split still owns its partial tiles: zeros, so the reduction can read every slot
        if (warp >= 2)
            for (int i = threadIdx.x - 64; i < kTaps * kTile / 4; i += 128) reinterpret_cast<float4*>(part)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else if (warp == 0) {
        if (lane == 0) {
            const uint32_t tx_bytes = kDzBytes + 2 * halo_rows * 128;
            for (int i = 0; i < n_iter; ++i) {
                const int s = i % kStages;
                mbar_wait(&empty_bar[s], ((i / kStages) & 1) ^ 1);
                const int chunk = c_beg + i;
                const int b = chunk / chunks_per_sample;
                const int l0 = (chunk % chunks_per_sample) * 64;
                uint8_t* a_dst = smem + s * kStageBytes;
                uint8_t* b_dst = a_dst + kDzBytes;
                mbar_expect_tx(&full_bar[s], tx_bytes);
                tma_load_3d(a_dst, mapP, &full_bar[s], 0, l0, b);
                tma_load_3d(a_dst + 8192, mapP, &full_bar[s], 64, l0, b);
                // positions outside [0, L) come back as zeros: the 'same' padding, and no leakage between samples
                tma_load_3d(b_dst, mapX, &full_bar[s], 0, l0 + shift0, b);
                tma_load_3d(b_dst + kRegion, mapX, &full_bar[s], 64, l0 + shift0, b);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(128, 128, 1, 1);
            for (int i = 0; i < n_iter; ++i) {
                const int s = i % kStages;
                mbar_wait(&full_bar[s], (i / kStages) & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
                const uint64_t da0 = make_smem_desc_sw128(a_addr, 8192, 1024);
#pragma unroll
                for (int j = 0; j < kTaps; ++j) {
                    const uint64_t db0 = desc_mn_shifted(a_addr + kDzBytes + j * d * 128, kRegion);
#pragma unroll
                    for (int k = 0; k < 4; ++k)      // 16 positions per MMA = 2048 B down the tile = +128 in the address field
                        umma_bf16(tmem_base + j * 128, da0 + (uint64_t)(k * 128), db0 + (uint64_t)(k * 128), idesc, (i == 0 && k == 0) ? 0u : 1u);
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(tmem_full_bar);
        }
    } else {
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        const int q = warp & 3;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
        for (int j = 0; j < kTaps; ++j) {
            float* wrow = part + (size_t)j * kTile + (size_t)(q * 32 + lane) * 128;
#pragma unroll 1
            for (int ch = 0; ch < 4; ++ch) {
                float acc[32];
                tmem_ld32(taddr + j * 128 + ch * 32, acc);
#pragma unroll
                for (int v = 0; v < 8; ++v)
                    reinterpret_cast<float4*>(wrow + ch * 32)[v] = make_float4(acc[4 * v], acc[4 * v + 1], acc[4 * v + 2], acc[4 * v + 3]);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// out[src][co, ci, t] = sum over splits.  One CTA per (co, src): 32 float4 lanes over ci x one thread row per tap read the
// partial tiles, twelve independent 16-byte loads in flight per thread (the kernel is a chain of L2 round trips); the
// [ci][t] row is assembled in shared memory and leaves as one contiguous run of 128 * ksize floats.
__global__ void __launch_bounds__(512) conv_wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ out0,
                                                                float* __restrict__ out1, int splits, int ksize) {
    extern __shared__ float row[];                    // [128][ksize]
    pdl_wait();
    const int m = blockIdx.x, src = blockIdx.y;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, nty = blockDim.x >> 5;
    const size_t total4 = (size_t)gridDim.y * ksize * kTile / 4;
    for (int t = ty; t < ksize; t += nty) {
        const float4* p = reinterpret_cast<const float4*>(ws + ((size_t)(src * ksize + t) * 128 + m) * 128) + tx;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a;
        int k = 0;
        for (; k + 12 <= splits; k += 12) {
            float4 v[12];
#pragma unroll
            for (int i = 0; i < 12; ++i) v[i] = p[(size_t)(k + i) * total4];
#pragma unroll
            for (int i = 0; i < 12; i += 2) {
                a.x += v[i].x; a.y += v[i].y; a.z += v[i].z; a.w += v[i].w;
                c.x += v[i + 1].x; c.y += v[i + 1].y; c.z += v[i + 1].z; c.w += v[i + 1].w;
            }
        }
        for (; k < splits; ++k) {
            const float4 v = p[(size_t)k * total4];
            a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
        }
        row[(tx * 4 + 0) * ksize + t] = a.x + c.x;
        row[(tx * 4 + 1) * ksize + t] = a.y + c.y;
        row[(tx * 4 + 2) * ksize + t] = a.z + c.z;
        row[(tx * 4 + 3) * ksize + t] = a.w + c.w;
    }
    __syncthreads();
    float* dst = (src == 0 ? out0 : out1) + (size_t)m * 128 * ksize;
    for (int i = threadIdx.x; i < 128 * ksize; i += blockDim.x) dst[i] = row[i];
}

}  // namespace

bool pb_tc_conv_wgrad3_ok(int B, int L, int C, int ksize, int dil, size_t ws_bytes) {
    if (!(pb_tc_enabled() && C == 128 && L % 64 == 0 && B >= 1 && ksize % kTaps == 0 && (ksize & 1))) return false;
    if (64 + (kTaps - 1) * dil > kRegion / 128 || dil < 1) return false;
    return ws_bytes >= (size_t)148 * kTaps * kTile * sizeof(float);
}

int pb_tc_conv_wgrad3(const void* dzn, const void* dzw, const void* x, float* dWn, float* dWw, int B, int L, int ksize, int dil,
                      void* ws, cudaStream_t st) {
    const int C = 128;
    CUtensorMap mP0, mP1, mX0, mX1;
    const uint64_t dims[3] = {(uint64_t)C, (uint64_t)L, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)C * 2, (uint64_t)L * C * 2};
    const uint32_t box_dz[3] = {64, 64, 1};
    const uint32_t box_x0[3] = {64, (uint32_t)(64 + (kTaps - 1) * 1), 1};
    const uint32_t box_x1[3] = {64, (uint32_t)(64 + (kTaps - 1) * dil), 1};
    int rc;
    if ((rc = pb_make_tmap_bf16(&mP0, dzn, 3, dims, strides, box_dz))) return rc;
    if ((rc = pb_make_tmap_bf16(&mP1, dzw, 3, dims, strides, box_dz))) return rc;
    if ((rc = pb_make_tmap_bf16(&mX0, x, 3, dims, strides, box_x0))) return rc;
    if ((rc = pb_make_tmap_bf16(&mX1, x, 3, dims, strides, box_x1))) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        PB_CUDA(cudaFuncSetAttribute(tc_conv_wgrad3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        attr_set = true;
    }
    ConvWgradParams wp;
    wp.L = L; wp.ksize = ksize; wp.half = ksize / 2; wp.dil[0] = 1; wp.dil[1] = dil;
    wp.chunks_total = B * (L / 64);
    wp.ws = (float*)ws;
    const int groups = ksize / kTaps;
    int splits = 148 / (2 * groups);
    if (splits < 1) splits = 1;
    if (splits > wp.chunks_total) splits = wp.chunks_total;
    wp.chunks_per_split = (wp.chunks_total + splits - 1) / splits;
    splits = (wp.chunks_total + wp.chunks_per_split - 1) / wp.chunks_per_split;
    PB_CUDA(pb_launch_pdl(tc_conv_wgrad3_kernel, dim3(groups, 2, splits), dim3(kThreads), kSmemBytes, st, mP0, mP1, mX0, mX1, wp));
    const int red_threads = 32 * (ksize < 16 ? ksize : 16);                 // one thread row per tap
    PB_CUDA(pb_launch_pdl(conv_wgrad_reduce_kernel, dim3(128, 2), dim3(red_threads), (size_t)128 * ksize * sizeof(float), st, (const float*)wp.ws,
                          dWn, dWw, splits, ksize));
    pb_count_launch(2);
    return pb_check_launch("tc_conv_wgrad3");
}
