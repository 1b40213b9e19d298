// Narrow + dilated Conv1d pair as ONE bf16 implicit-GEMM kernel on tcgen05 tensor cores (sm_100a).
//
//   forward : zn = conv(x, Wn, d=1) + bn ; zw = conv(x, Ww, d=dil) + bw ;
//             u  = x + gelu(zn) + gelu(zw) + g2l[b,:] ; stats[b] += (sum u, sum u^2)      (modules.py:205-212)
//   dgrad   : dx = du + conv^T(dzn, Wn, d=1) + conv^T(dzw, Ww, d=dil)
//
// GEMM view per 128-position tile of one sample: D[128, C] = sum over taps t, channel chunks kc of
//   A_t[128, 64] (rows l0 + shift(t) .. +127 of the input, fetched by a 3-D TMA box whose out-of-range rows are
//   zero-filled = 'same' padding, never crossing into the neighbouring sample)  x  W_t[C, 64]^T.
// Both convolutions accumulate into their own 128 x C fp32 accumulator in TMEM (2*C of the 512 columns);
// the epilogue reads both, applies bias / GELU / residual / broadcast-add, writes bf16 and the LayerNorm sums.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
// warps 2..5 = epilogue (warp w owns TMEM lanes 32*(w%4)..+31).  3-stage smem ring, 32 KB per stage, so two
// CTAs share an SM and one CTA's epilogue overlaps the other's main loop.
#include "tc_common.cuh"
#include "tc_api.cuh"

namespace {

constexpr int kC = 128;                // channels handled by this kernel (N of the MMA, and 2 K-chunks of 64)
constexpr int kBM = 128;               // sequence positions per CTA
constexpr int kStages = 3;
constexpr int kStageBytes = 2 * kBM * 64 * 2;     // A tile + B tile, 16 KB each
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /* alignment slack */ + 256 /* barriers */;
constexpr int kThreads = 192;

struct ConvEpiParams {
    // forward
    const bf16* x; const float* bn; const float* bw; const float* g2l;
    bf16* zn; bf16* zw; bf16* u; double* stats;
    // dgrad
    const bf16* du; bf16* dx;
};

template <bool kFwd>
__global__ void __launch_bounds__(kThreads, 2)
tc_conv_pair_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
                    const __grid_constant__ CUtensorMap mapB0, const __grid_constant__ CUtensorMap mapB1,
                    ConvEpiParams ep, int L, int ksize, int dil) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + kStages * kStageBytes);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_per_sample = L / kBM;
    const int b = blockIdx.x / tiles_per_sample;
    const int l0 = (blockIdx.x % tiles_per_sample) * kBM;
    const int half = ksize / 2;
    const int kchunks = kC / 64;
    const int n_iter = 2 * ksize * kchunks;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA0); tma_prefetch_desc(&mapA1); tma_prefetch_desc(&mapB0); tma_prefetch_desc(&mapB1);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<2 * kC>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();          // everything above overlapped the previous kernel's tail; from here on we touch its outputs

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            for (int i = 0; i < n_iter; ++i) {
                const int s = i % kStages;
                const uint32_t ph = (i / kStages) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);
                const int conv = i / (ksize * kchunks);
                const int t = (i / kchunks) % ksize;
                const int kc = i % kchunks;
                const int d = conv == 0 ? 1 : dil;
                const int shift = (kFwd ? (t - half) : (half - t)) * d;
                uint8_t* a_dst = smem + s * kStageBytes;
                uint8_t* b_dst = a_dst + kBM * 64 * 2;
                mbar_expect_tx(&full_bar[s], kStageBytes);
                tma_load_3d(a_dst, conv == 0 ? &mapA0 : &mapA1, &full_bar[s], kc * 64, l0 + shift, b);
                tma_load_2d(b_dst, conv == 0 ? &mapB0 : &mapB1, &full_bar[s], kc * 64, t * kC);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(kBM, kC, 0, 0);
            for (int i = 0; i < n_iter; ++i) {
                const int s = i % kStages;
                const uint32_t ph = (i / kStages) & 1;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                const int conv = i / (ksize * kchunks);
                const int first = (i % (ksize * kchunks)) == 0;
                const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
                const uint64_t da0 = make_smem_desc_sw128(a_addr, 16, 1024);
                const uint64_t db0 = make_smem_desc_sw128(a_addr + kBM * 64 * 2, 16, 1024);
#pragma unroll
                for (int k = 0; k < 4; ++k)       // K step of 16 elements = +32 B = +2 in the descriptor's address field
                    umma_bf16(tmem_base + conv * kC, da0 + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), idesc, (first && k == 0) ? 0u : 1u);
                umma_commit(&empty_bar[s]);          // frees the smem stage when these MMAs retire
            }
            umma_commit(tmem_full_bar);              // accumulators complete
        }
    } else {
        // ===== epilogue: TMEM -> registers -> global =====
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const size_t grow = ((size_t)b * L + l0 + row) * kC;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        float s_sum = 0.f, s_sq = 0.f;
#pragma unroll 1
        for (int ch = 0; ch < kC / 32; ++ch) {
            float an[32], aw[32];
            tmem_ld32(taddr + ch * 32, an);
            tmem_ld32(taddr + kC + ch * 32, aw);
            const int c0 = ch * 32;
            if (kFwd) {
                uint4 xv[4];
#pragma unroll
                for (int v = 0; v < 4; ++v) xv[v] = *reinterpret_cast<const uint4*>(ep.x + grow + c0 + v * 8);
                const bf16* xb = reinterpret_cast<const bf16*>(xv);
                __align__(16) bf16 on[32], ow[32], ou[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int c = c0 + j;
                    const bf16 zn = __float2bfloat16_rn(an[j] + __ldg(ep.bn + c));
                    const bf16 zw = __float2bfloat16_rn(aw[j] + __ldg(ep.bw + c));
                    on[j] = zn; ow[j] = zw;
                    float gn, gw;
                    gelu_fast2(__bfloat162float(zn), __bfloat162float(zw), gn, gw);
                    float v = __bfloat162float(xb[j]) + gn;
                    v += gw;
                    v += __ldg(ep.g2l + (size_t)b * kC + c);
                    const bf16 ub = __float2bfloat16_rn(v);
                    ou[j] = ub;
                    const float ur = __bfloat162float(ub);
                    s_sum += ur; s_sq += ur * ur;
                }
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    *reinterpret_cast<uint4*>(ep.zn + grow + c0 + v * 8) = reinterpret_cast<const uint4*>(on)[v];
                    *reinterpret_cast<uint4*>(ep.zw + grow + c0 + v * 8) = reinterpret_cast<const uint4*>(ow)[v];
                    *reinterpret_cast<uint4*>(ep.u + grow + c0 + v * 8) = reinterpret_cast<const uint4*>(ou)[v];
                }
            } else {
                uint4 dv[4];
#pragma unroll
                for (int v = 0; v < 4; ++v) dv[v] = *reinterpret_cast<const uint4*>(ep.du + grow + c0 + v * 8);
                const bf16* db_ = reinterpret_cast<const bf16*>(dv);
                __align__(16) bf16 od[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) od[j] = __float2bfloat16_rn(__bfloat162float(db_[j]) + an[j] + aw[j]);
#pragma unroll
                for (int v = 0; v < 4; ++v)
                    *reinterpret_cast<uint4*>(ep.dx + grow + c0 + v * 8) = reinterpret_cast<const uint4*>(od)[v];
            }
        }
        if (kFwd) {
            const double ds = warp_sum_d((double)s_sum), dq = warp_sum_d((double)s_sq);
            if (lane == 0) { atomicAdd(&ep.stats[2 * b], ds); atomicAdd(&ep.stats[2 * b + 1], dq); }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<2 * kC>(tmem_base);
    }
}

int make_act_map(CUtensorMap* m, const void* base, int B, int L, int C, int box_rows) {
    const uint64_t dims[3] = {(uint64_t)C, (uint64_t)L, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)C * 2, (uint64_t)L * C * 2};
    const uint32_t box[3] = {64, (uint32_t)box_rows, 1};
    return pb_make_tmap_bf16(m, base, 3, dims, strides, box);
}
int make_mat_map(CUtensorMap* m, const void* base, long rows, int cols, int box_rows) {
    const uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
    const uint64_t strides[1] = {(uint64_t)cols * 2};
    const uint32_t box[2] = {64, (uint32_t)box_rows};
    return pb_make_tmap_bf16(m, base, 2, dims, strides, box);
}

template <bool kFwd>
int launch_conv(const void* a0, const void* a1, const void* w0, const void* w1, const ConvEpiParams& ep, int B, int L, int ksize,
                int dil, cudaStream_t st) {
    CUtensorMap mA0, mA1, mB0, mB1;
    int rc;
    if ((rc = make_act_map(&mA0, a0, B, L, kC, kBM))) return rc;
    if ((rc = make_act_map(&mA1, a1, B, L, kC, kBM))) return rc;
    if ((rc = make_mat_map(&mB0, w0, (long)ksize * kC, kC, kC))) return rc;
    if ((rc = make_mat_map(&mB1, w1, (long)ksize * kC, kC, kC))) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        PB_CUDA(cudaFuncSetAttribute(tc_conv_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        PB_CUDA(cudaFuncSetAttribute(tc_conv_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        attr_set = true;
    }
    PB_CUDA(pb_launch_pdl(tc_conv_pair_kernel<kFwd>, dim3(B * (L / kBM)), dim3(kThreads), kSmemBytes, st, mA0, mA1, mB0, mB1, ep, L, ksize, dil));
    pb_count_launch(1);
    return pb_check_launch(kFwd ? "tc_conv_pair_fwd" : "tc_conv_pair_dgrad");
}

}  // namespace

bool pb_tc_conv_ok(int B, int L, int C, int ksize, int dil) {
    if (!pb_tc_enabled() || !(ksize >= 1 && (ksize & 1) && dil >= 1 && B >= 1)) return false;
    if (C == kC && L % kBM == 0) return true;                                  // this file's kernel (C = 128)
    return pb_tc_mode() != 4 && pb_tc_conv2_ok(B, L, C, ksize, dil);           // C = 256: the halo kernel only
}

int pb_tc_conv_pair_fwd(const void* x, const void* wn, const void* ww, const float* bn, const float* bw, const float* g2l,
                        void* zn, void* zw, void* u, double* stats, int B, int L, int C, int ksize, int dil, cudaStream_t st) {
    const int mode = pb_tc_mode();
    if (mode != 4 && pb_tc_conv2_ok(B, L, C, ksize, dil))       // 256-position halo kernel unless the first generation is forced
        return pb_tc_conv2_fwd(x, wn, ww, bn, bw, g2l, zn, zw, u, stats, B, L, C, ksize, dil, 0, st);
    ConvEpiParams ep{};
    ep.x = (const bf16*)x; ep.bn = bn; ep.bw = bw; ep.g2l = g2l;
    ep.zn = (bf16*)zn; ep.zw = (bf16*)zw; ep.u = (bf16*)u; ep.stats = stats;
    return launch_conv<true>(x, x, wn, ww, ep, B, L, ksize, dil, st);
}

int pb_tc_conv_pair_dgrad(const void* du, const void* dzn, const void* dzw, const void* wn_bwd, const void* ww_bwd, void* dx,
                          int B, int L, int C, int ksize, int dil, cudaStream_t st) {
    const int mode = pb_tc_mode();
    if (mode != 4 && pb_tc_conv2_ok(B, L, C, ksize, dil))
        return pb_tc_conv2_dgrad(du, dzn, dzw, wn_bwd, ww_bwd, dx, B, L, C, ksize, dil, 0, st);
    ConvEpiParams ep{};
    ep.du = (const bf16*)du; ep.dx = (bf16*)dx;
    return launch_conv<false>(dzn, dzw, wn_bwd, ww_bwd, ep, B, L, ksize, dil, st);
}
