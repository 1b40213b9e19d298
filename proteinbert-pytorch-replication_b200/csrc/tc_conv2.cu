// Conv pair, second-generation tcgen05 kernel: 256 sequence positions per CTA, input halo resident in shared memory.
//
// What changed against tc_conv.cu (which fetched one 128 x 64 activation tile per tap, 18 taps x 2 chunks per
// tile, and was bound by L2 -> shared-memory traffic at ~1.15 MB per 128 positions):
//   * the CTA loads the input rows it needs ONCE: positions [l0 - margin, l0 + 256 + margin) x 128 channels
//     (74 KB for the dilated conv, margin = 4 * dilation) and every tap's A operand is the same tile read through
//     a UMMA descriptor whose start address is moved down by (tap offset) rows of 128 bytes;
//   * each 16 KB weight tile that arrives in the ring feeds 8 MMAs (two 128-position sub-tiles x 4 k-steps)
//     instead of 4, so the weight stream per position is halved;
//   * four accumulators (2 sub-tiles x 2 convolutions) fill the 512 TMEM columns; eight epilogue warps;
//   * forward: the GELU epilogue is 60 % as long as the MMA phase (cycle counters: 19 k vs 31 k cycles), so it is cut in
//     two: the narrow convolution's half (zn out, x + gelu(zn) + g2l written back into its TMEM columns with
//     tcgen05.st) runs while the dilated convolution's MMAs are still being issued; only the second half is exposed.
// L2 -> smem traffic drops to ~0.65 MB per 256 positions (3.5x less per position).
//
// Round 2: operand roles swapped.  With positions as the MMA's M (two 128 x 128 x 16 MMAs per weight k-step) the tensor pipe
// waited on shared memory: each MMA read 4 KB of activations + 4 KB of weights, 128 B per cycle at the 64-cycle rate, and
// issued at ~107 cycles (profiles/r01_conv2_fwd_ncu.md).  Now the WEIGHT tile is the A operand (M = 128 output channels)
// and all 256 positions of the CTA are the B operand (N = 256, the same row-shifted halo descriptor): one 128 x 256 x 16
// MMA per k-step reads 4 KB + 8 KB for twice the work -- 25 % less shared-memory traffic per FLOP, half the instructions.
// The accumulators come out transposed (TMEM lane = output channel, column = position), which makes bias and the global
// broadcast per-lane constants; the epilogue transposes back on its way to the staging tiles (one shuffle per two values:
// neighbouring lanes = neighbouring channels form the 32-bit words of a [position][channel] row).
#include <stdlib.h>
#include "tc_common.cuh"
#include "tc_api.cuh"

namespace {

constexpr int kBTile = 128 * 64 * 2;             // one weight tile [128 out-ch x 64 in-ch] bf16 = 16 KB
// Channel count C is a template parameter: 128 (the default model) or 256 (BASELINE config 5).  The 512 TMEM columns hold
// [C/128 output-channel tiles] x [positions] x [2 convolutions] accumulators, so a CTA takes 256 positions at C = 128 and
// 128 positions at C = 256; either way the epilogue sees two 128-column halves per accumulator.
template <int kC> struct ConvCfg {
    static constexpr int kChunks = kC / 64;      // 64-channel K chunks (one halo buffer each)
    static constexpr int kCoTiles = kC / 128;    // 128-channel output tiles (MMA M)
    static constexpr int kN = kC == 128 ? 256 : 128;      // positions per CTA (MMA N)
};
constexpr int kThreads = 320;                    // warp 0 TMA, warp 1 MMA + TMEM, warps 2..9 epilogue

struct Conv2Params {
    const bf16* x; const float* bn; const float* bw; const float* g2l;
    bf16* zn; bf16* zw; bf16* u; double* stats;
    const bf16* du; bf16* dx;
    int L, ksize, dil;
    int hr[2];            // halo rows of input 0 / 1
    int a_bytes[2];       // shared-memory bytes of one (input, channel chunk) halo tile, multiple of 1024
    int both_halos;       // data gradient at C = 256: both inputs' halos are resident from the start (no refill)
    unsigned long long* dbg;   // optional cycle counters of CTA 0 (pb_debug_buffer)
    int debug;            // experiment switches (PB_CONV2_DEBUG): 1 = align tap offsets to 8 rows (wrong results), 2 = skip epilogue math
    int base_offset_mode; // 0 (measured on B200): the 128-byte swizzle is a function of the absolute shared-memory
                          // address, so a descriptor that starts on any row of a TMA-written tile needs base_offset = 0;
                          // 1 = (addr >> 7) & 7, kept only to document the convention that does NOT work
};

__device__ __forceinline__ uint64_t desc_sw128_bo(uint32_t saddr, uint32_t sbo_bytes, int mode) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;                                  // LBO (unused for swizzled K-major)
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;
    if (mode) d |= (uint64_t)((saddr >> 7) & 0x7u) << 49;    // start row inside the 8-row swizzle atom
    d |= (uint64_t)2 << 61;
    return d;
}

template <bool kFwd, int kStages, int kC>
__global__ void __launch_bounds__(kThreads, 1)
tc_conv2_kernel(const __grid_constant__ CUtensorMap mapA0m, const __grid_constant__ CUtensorMap mapA0t,
                const __grid_constant__ CUtensorMap mapA1m, const __grid_constant__ CUtensorMap mapA1t,
                const __grid_constant__ CUtensorMap mapB0, const __grid_constant__ CUtensorMap mapB1,
                const __grid_constant__ CUtensorMap mapS0, const __grid_constant__ CUtensorMap mapS1,
                const __grid_constant__ CUtensorMap mapS2, Conv2Params p) {
    constexpr int kChunks = ConvCfg<kC>::kChunks, kCoTiles = ConvCfg<kC>::kCoTiles, kRows = ConvCfg<kC>::kN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    // layout: halo buffers (one per 64-channel chunk, sized for the larger input; two sets when both inputs stay
    // resident) | weight ring | barriers.
    // forward: the buffers hold x (wide margin) for the whole kernel.  dgrad at C = 128: they hold dzn first; as soon as
    // the narrow convolution has consumed chunk kc, the same buffer is refilled with dzw for the dilated one.  dgrad at
    // C = 256: dzn and dzw each have their own buffers (the 128-position halos are small enough).
    const int a_buf = p.a_bytes[0] > p.a_bytes[1] ? p.a_bytes[0] : p.a_bytes[1];
    const int n_abuf = p.both_halos ? 2 * kChunks : kChunks;
    uint8_t* ring = smem + n_abuf * a_buf;
    // forward: the first half of the epilogue (narrow convolution) runs while the ring still streams the dilated
    // convolution's weights, so the staging tiles get their own area (one 3 x 2 KB buffer per epilogue warp) behind the
    // ring; dgrad stages inside the halo area, dead by then.
    uint8_t* stage_base = kFwd ? ring + kStages * kBTile : smem;
    uint64_t* full_bar = (uint64_t*)(kFwd ? stage_base + 8 * 3 * 2048 : ring + kStages * kBTile);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* a_full = empty_bar + kStages;          // [2 inputs][kChunks]
    uint64_t* a_free = a_full + 2 * kChunks;         // [kChunks]: narrow conv done with the buffer (dgrad, refill mode)
    uint64_t* tmem_full_bar = a_free + kChunks;      // [2]: forward: narrow conv's accumulators / all; dgrad uses [1] only
    uint32_t* tmem_slot = (uint32_t*)(tmem_full_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_per_sample = p.L / kRows;
    const int b = blockIdx.x / tiles_per_sample;
    const int l0 = (blockIdx.x % tiles_per_sample) * kRows;
    const int half = p.ksize / 2;
    auto abuf_of = [&](int in, int kc) { return smem + (size_t)((p.both_halos ? in * kChunks : 0) + kc) * a_buf; };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA0m); tma_prefetch_desc(&mapA1m); tma_prefetch_desc(&mapB0); tma_prefetch_desc(&mapB1);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int i = 0; i < 2 * kChunks; ++i) mbar_init(&a_full[i], 1);
        for (int i = 0; i < kChunks; ++i) mbar_init(&a_free[i], 1);
        mbar_init(&tmem_full_bar[0], 1); mbar_init(&tmem_full_bar[1], 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    // iteration i -> (conv, output tile, channel chunk, tap): chunk-major inside an output tile so that at C = 128 a halo
    // buffer is released as early as possible
    if (warp == 0) {
        if (lane == 0) {
            auto load_halo = [&](int in, int kc) {
                const int margin = half * (in == 0 ? 1 : p.dil);
                const int tail_rows = p.hr[in] - kRows;
                uint64_t* bar = &a_full[in * kChunks + kc];
                uint8_t* dst = abuf_of(in, kc);
                mbar_expect_tx(bar, p.hr[in] * 128);
                tma_load_3d(dst, in == 0 ? &mapA0m : &mapA1m, bar, kc * 64, l0 - margin, b);
                if (tail_rows > 0)
                    tma_load_3d(dst + kRows * 128, in == 0 ? &mapA0t : &mapA1t, bar, kc * 64, l0 - margin + kRows, b);
            };
            const int in_first = kFwd ? 1 : 0;
            for (int kc = 0; kc < kChunks; ++kc) load_halo(in_first, kc);
            if (!kFwd && p.both_halos)
                for (int kc = 0; kc < kChunks; ++kc) load_halo(1, kc);
            int s = 0;
            uint32_t ph = 1;                                 // parity to wait for on empty_bar[s]: flips each time the ring wraps
            for (int conv = 0; conv < 2; ++conv)
                for (int cot = 0; cot < kCoTiles; ++cot)
                    for (int kc = 0; kc < kChunks; ++kc)
                        for (int t = 0; t < p.ksize; ++t) {
                            if (!kFwd && !p.both_halos && conv == 1 && t == 0) {   // refill this chunk's buffer with the dilated conv's input
                                mbar_wait(&a_free[kc], 0);
                                load_halo(1, kc);
                            }
                            mbar_wait(&empty_bar[s], ph);
                            mbar_expect_tx(&full_bar[s], kBTile);
                            tma_load_2d(ring + s * kBTile, conv == 0 ? &mapB0 : &mapB1, &full_bar[s], kc * 64, t * kC + cot * 128);
                            if (++s == kStages) { s = 0; ph ^= 1; }
                        }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(128, kRows, 0, 0);       // M = 128 output channels, N = positions of the CTA
#ifdef PB_CONV2_PROBE
            long long t_wait_full = 0, t_wait_a = 0;
            const long long t_begin = clock64();
#endif
            int s = 0;
            uint32_t ph = 0;                                 // parity to wait for on full_bar[s]
            for (int conv = 0; conv < 2; ++conv) {
                const int in = kFwd ? 1 : conv;
                const int d = conv == 0 ? 1 : p.dil;
                for (int cot = 0; cot < kCoTiles; ++cot) {
                    // forward: one accumulator set per convolution (the narrow one is drained while the dilated one is still
                    // being formed); data gradient: both convolutions add into the same accumulators.  Output tile `cot` owns
                    // columns [cot*128, +128) of its set at C = 256; at C = 128 the single tile spans all 256.
                    const uint32_t tm = tmem_base + (kFwd ? conv * 256 : 0) + cot * 128;
                    for (int kc = 0; kc < kChunks; ++kc) {
                        const uint32_t abase = smem_u32(smem + (size_t)((p.both_halos ? in * kChunks : 0) + kc) * a_buf);
                        for (int t = 0; t < p.ksize; ++t) {
                            // row of the halo tile at which this tap's window starts
                            int row_off = kFwd ? half * (p.dil - d) + t * d : (2 * half - t) * d;
                            if (p.debug & 1) row_off &= ~7;
#ifdef PB_CONV2_PROBE
                            long long c0 = clock64();
#endif
                            if (t == 0 && cot == 0 && (conv == 0 || !kFwd)) mbar_wait(&a_full[in * kChunks + kc], 0);
#ifdef PB_CONV2_PROBE
                            long long c1 = clock64();
#endif
                            mbar_wait(&full_bar[s], ph);
#ifdef PB_CONV2_PROBE
                            t_wait_a += c1 - c0; t_wait_full += clock64() - c1;
#endif
                            tc_fence_after();
                            // one descriptor pair per stage; the per-MMA variants differ only in the 16-byte-granular start
                            // address field (bits 0..13), so they are formed by adding constants -- the issuing thread is the
                            // bottleneck (no divisions in this loop either)
                            const uint64_t dact = desc_sw128_bo(abase + (uint32_t)row_off * 128u, 1024, p.base_offset_mode);
                            const uint64_t dwgt = make_smem_desc_sw128(smem_u32(ring + s * kBTile), 16, 1024);
                            const uint32_t acc0 = kFwd ? ((t == 0 && kc == 0) ? 0u : 1u) : ((conv == 0 && t == 0 && kc == 0) ? 0u : 1u);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                umma_bf16(tm, dwgt + (uint64_t)(k * 2), dact + (uint64_t)(k * 2), idesc, k == 0 ? acc0 : 1u);
                            umma_commit(&empty_bar[s]);
                            if (!kFwd && !p.both_halos && conv == 0 && t == p.ksize - 1) umma_commit(&a_free[kc]);   // buffer kc may be refilled
                            if (++s == kStages) { s = 0; ph ^= 1; }
                        }
                    }
                }
                if (kFwd && conv == 0) umma_commit(&tmem_full_bar[0]);          // narrow conv complete
            }
            umma_commit(&tmem_full_bar[1]);
#ifdef PB_CONV2_PROBE
            if (p.dbg && blockIdx.x == 0) {
                const long long t_issue_end = clock64();
                mbar_wait(&tmem_full_bar[1], 0);
                const long long t_done = clock64();
                p.dbg[0] = t_issue_end - t_begin; p.dbg[1] = t_wait_full; p.dbg[2] = t_wait_a; p.dbg[3] = t_done - t_begin;
            }
#endif
        }
    } else {
        // ===== epilogue: warps 2..9.  TMEM lane = output channel: quarter q = warp % 4 -> channels 32q .. 32q+31 (one per
        //       lane); position half = (warp-2)/4 -> columns 128*half .. +127 of each accumulator, in chunks of 32 =====
#ifdef PB_CONV2_PROBE
        const long long e0 = clock64();
        long long e1 = e0;
#endif
        // C = 128: the warp's half selects positions 128*hsel .. +127 of the single output tile; C = 256: it selects the output
        // tile (channels 128*hsel ..) and the warp covers all 128 positions.  Either way: TMEM columns 128*hsel .. +127.
        const int hsel = (warp - 2) >> 2, q = warp & 3;
        const int hpos = kC == 128 ? hsel : 0;                          // position half
        const int co = (kC == 128 ? 0 : hsel * 128) + q * 32 + lane;    // this thread's channel
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + hsel * 128;
        // Output tiles leave through TMA: each warp stages a [32 positions x 32 channels] bf16 tile per tensor in shared
        // memory (64-byte swizzle) and one lane issues the bulk tensor store.
        uint8_t* stage = stage_base + (warp - 2) * (kFwd ? 3 * 2048 : 2 * 2048);
        const int lrow = l0 + hpos * 128;                               // first position of this warp's half
        // transposing store: v[j] = this lane's channel at position j goes to row j of a [32 positions x 32 channels] tile
        // (64-byte rows, 64-byte swizzle: 16-byte chunk index ^ ((row >> 1) & 3)).  One 2-byte store per value: the 32 lanes of a
        // warp write the 64 contiguous bytes of one row (two lanes per 32-bit word, one wavefront).  The swizzle term takes four
        // values, so the four lane-dependent base offsets are formed once and every store address is base + constant.
        // (The first version packed pairs through a shuffle to store 32-bit words: 16 stores instead of 32 per tile, but 48 more
        // ALU instructions, and this epilogue is bound by instruction issue.)
        uint32_t sw_off[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) sw_off[k] = ((((uint32_t)lane >> 3) ^ (uint32_t)k) << 4) + ((uint32_t)lane & 7u) * 2u;
        auto stage_transposed = [&](uint8_t* tile, const bf16 (&v)[32]) {
#pragma unroll
            for (int j = 0; j < 32; ++j) *reinterpret_cast<bf16*>(tile + j * 64 + sw_off[(j >> 1) & 3]) = v[j];
        };
        float s_sum = 0.f, s_sq = 0.f;
        if (kFwd) {
            // ---- first half, under the MMAs of the dilated convolution: zn = acc + bn (stored), and the partial sum
            //      x + gelu(zn) + g2l goes back into the accumulator's TMEM columns (tcgen05.st) for the second half.
            //      x comes from the halo tile that feeds the MMAs.
            const float bnl = __ldg(p.bn + co), gl = __ldg(p.g2l + (size_t)b * kC + co);
            // x[position][co] sits in the 128-byte-swizzled halo tile: 16-byte chunk index ^ (row & 7).  The chunk's first halo
            // row is a multiple of 4 ((2 * half * dil) % 8 == 0, positions in steps of 32), so (row0 + j) & 7 == (j & 7) ^ (row0 & 7):
            // the per-position part of the swizzle is a compile-time constant
            const int xsw = (((co & 63) >> 3) ^ ((half * p.dil) & 7)) << 4;
            const uint8_t* xbase = smem + (size_t)(co >> 6) * a_buf + (co & 7) * 2;
            uint32_t xoff[8];                       // swizzled chunk offset for positions j with (j & 7) == k
#pragma unroll
            for (int k = 0; k < 8; ++k) xoff[k] = (uint32_t)(xsw ^ (k << 4));
            mbar_wait(&tmem_full_bar[0], 0);
#ifdef PB_CONV2_PROBE
            e1 = clock64();
#endif
            tc_fence_after();
#pragma unroll 1
            for (int ch = 0; ch < ((p.debug & 2) ? 0 : 4); ++ch) {
                float an[32];
                tmem_ld32(taddr + ch * 32, an);
                uint8_t* sbuf = stage + (ch & 1) * 2048;
                const uint8_t* xrow = xbase + (size_t)(half * p.dil + hpos * 128 + ch * 32) * 128;     // halo row of the chunk's first position
                __align__(16) bf16 on[32];
                float gz[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const bf16 zn = __float2bfloat16_rn(an[j] + bnl);
                    on[j] = zn;
                    gz[j] = __bfloat162float(zn);
                }
                gelu_fast_n(gz);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const bf16 xv = *reinterpret_cast<const bf16*>(xrow + j * 128 + xoff[j & 7]);
                    an[j] = (__bfloat162float(xv) + gz[j]) + gl;
                }
                tmem_st32(taddr + ch * 32, an);
                if (ch >= 2) {                                           // the buffer's previous store must have read it
                    if (lane == 0) tma_store_wait_read<1>();
                    __syncwarp();
                }
                stage_transposed(sbuf, on);
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) { tma_store_3d(&mapS0, sbuf, co - lane, lrow + ch * 32, b); tma_store_commit(); }
            }
            // ---- second half, after the last MMA: zw = acc + bw, u = partial + gelu(zw), LayerNorm sums
            const float bwl = __ldg(p.bw + co);
            mbar_wait(&tmem_full_bar[1], 0);
            tc_fence_after();
#pragma unroll 1
            for (int ch = 0; ch < ((p.debug & 2) ? 0 : 4); ++ch) {
                float pa[32], aw[32];
                tmem_ld32(taddr + ch * 32, pa);
                tmem_ld32(taddr + 256 + ch * 32, aw);
                __align__(16) bf16 ow[32], ou[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const bf16 zw = __float2bfloat16_rn(aw[j] + bwl);
                    ow[j] = zw;
                    aw[j] = __bfloat162float(zw);
                }
                gelu_fast_n(aw);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const bf16 ub = __float2bfloat16_rn(pa[j] + aw[j]);
                    ou[j] = ub;
                    const float ur = __bfloat162float(ub);
                    s_sum += ur; s_sq += ur * ur;
                }
                if (lane == 0) tma_store_wait_read<0>();                 // one pair of staging tiles: the previous stores have read them
                __syncwarp();
                stage_transposed(stage, ow);
                stage_transposed(stage + 2048, ou);
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    tma_store_3d(&mapS1, stage, co - lane, lrow + ch * 32, b);
                    tma_store_3d(&mapS2, stage + 2048, co - lane, lrow + ch * 32, b);
                    tma_store_commit();
                }
            }
        } else {
            mbar_wait(&tmem_full_bar[1], 0);
#ifdef PB_CONV2_PROBE
            e1 = clock64();
#endif
            tc_fence_after();
#pragma unroll 1
            for (int ch = 0; ch < ((p.debug & 2) ? 0 : 4); ++ch) {
                // du[position][channel] for this lane's channel: 32 two-byte loads, 64 contiguous bytes per warp instruction
                const bf16* dup = p.du + ((size_t)b * p.L + lrow + ch * 32) * kC + co;
                bf16 dv[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) dv[j] = dup[(size_t)j * kC];
                float an[32];
                tmem_ld32(taddr + ch * 32, an);
                uint8_t* sbuf = stage + (ch & 1) * 2048;
                if (ch >= 2) {                                           // the buffer's previous store must have read it
                    if (lane == 0) tma_store_wait_read<1>();
                    __syncwarp();
                }
                __align__(16) bf16 od[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) od[j] = __float2bfloat16_rn(__bfloat162float(dv[j]) + an[j]);
                stage_transposed(sbuf, od);
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) { tma_store_3d(&mapS0, sbuf, co - lane, lrow + ch * 32, b); tma_store_commit(); }
            }
        }
        if (lane == 0) tma_store_wait<0>();
        if (kFwd) {
            const double ds = warp_sum_d((double)s_sum), dq = warp_sum_d((double)s_sq);
            if (lane == 0) { atomicAdd(&p.stats[2 * b], ds); atomicAdd(&p.stats[2 * b + 1], dq); }
        }
        tc_fence_before();
#ifdef PB_CONV2_PROBE
        if (p.dbg && blockIdx.x == 0 && threadIdx.x == 64) { p.dbg[4] = e1 - e0; p.dbg[5] = clock64() - e1; }
#endif
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

int act_map(CUtensorMap* m, const void* base, int B, int L, int C, int box_rows) {
    const uint64_t dims[3] = {(uint64_t)C, (uint64_t)L, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)C * 2, (uint64_t)L * C * 2};
    const uint32_t box[3] = {64, (uint32_t)box_rows, 1};
    return pb_make_tmap_bf16(m, base, 3, dims, strides, box);
}
int w_map(CUtensorMap* m, const void* base, int C, int ksize) {          // [k][Co][Ci] pack: rows = tap*Co + co, box = 128 x 64
    const uint64_t dims[2] = {(uint64_t)C, (uint64_t)ksize * C};
    const uint64_t strides[1] = {(uint64_t)C * 2};
    const uint32_t box[2] = {64, 128};
    return pb_make_tmap_bf16(m, base, 2, dims, strides, box);
}

template <bool kFwd, int kStages, int kC>
int launch2(const void* a0, const void* a1, const void* w0, const void* w1, Conv2Params p, int B, int L, cudaStream_t st) {
    constexpr int kRows = ConvCfg<kC>::kN, kChunks = ConvCfg<kC>::kChunks;
    const int half = p.ksize / 2;
    static const int debug_env = [] { const char* e = getenv("PB_CONV2_DEBUG"); return e ? atoi(e) : 0; }();   // read once
    p.debug = debug_env;
    p.dbg = (unsigned long long*)pb_debug_ptr();
    p.hr[0] = kRows + 2 * half;
    p.hr[1] = kRows + 2 * half * p.dil;
    for (int i = 0; i < 2; ++i) p.a_bytes[i] = (p.hr[i] * 128 + 1023) / 1024 * 1024;
    p.both_halos = (!kFwd && kC == 256) ? 1 : 0;
    CUtensorMap mA0m, mA0t, mA1m, mA1t, mB0, mB1, mS0, mS1, mS2;
    int rc;
    {   // store maps: 32 positions x 32 channels per box, 64-byte swizzle
        const uint64_t dims[3] = {(uint64_t)kC, (uint64_t)L, (uint64_t)B};
        const uint64_t strides[2] = {(uint64_t)kC * 2, (uint64_t)L * kC * 2};
        const uint32_t box[3] = {32, 32, 1};
        void* o0 = kFwd ? (void*)p.zn : (void*)p.dx;
        void* o1 = kFwd ? (void*)p.zw : (void*)p.dx;
        void* o2 = kFwd ? (void*)p.u : (void*)p.dx;
        if ((rc = pb_make_tmap_bf16(&mS0, o0, 3, dims, strides, box, 64))) return rc;
        if ((rc = pb_make_tmap_bf16(&mS1, o1, 3, dims, strides, box, 64))) return rc;
        if ((rc = pb_make_tmap_bf16(&mS2, o2, 3, dims, strides, box, 64))) return rc;
    }
    if ((rc = act_map(&mA0m, a0, B, L, kC, kRows))) return rc;
    if ((rc = act_map(&mA0t, a0, B, L, kC, p.hr[0] - kRows))) return rc;
    if ((rc = act_map(&mA1m, a1, B, L, kC, kRows))) return rc;
    if ((rc = act_map(&mA1t, a1, B, L, kC, p.hr[1] - kRows))) return rc;
    if ((rc = w_map(&mB0, w0, kC, p.ksize))) return rc;
    if ((rc = w_map(&mB1, w1, kC, p.ksize))) return rc;
    const size_t a_buf = p.a_bytes[0] > p.a_bytes[1] ? p.a_bytes[0] : p.a_bytes[1];
    const size_t smem = (size_t)(p.both_halos ? 2 : 1) * kChunks * a_buf + (size_t)kStages * kBTile + (kFwd ? 8 * 3 * 2048 : 0) + 1024 + 256;
    PB_REQUIRE(smem <= 232448, PB_ERR_UNSUPPORTED, "conv2: %zu bytes of shared memory needed", smem);
    static size_t smem_set = 0;
    if (smem_set != smem) {
        PB_CUDA(cudaFuncSetAttribute(tc_conv2_kernel<kFwd, kStages, kC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    PB_CUDA(pb_launch_pdl(tc_conv2_kernel<kFwd, kStages, kC>, dim3(B * (L / kRows)), dim3(kThreads), smem, st, mA0m, mA0t, mA1m, mA1t,
                          mB0, mB1, mS0, mS1, mS2, p));
    pb_count_launch(1);
    return pb_check_launch(kFwd ? "tc_conv2_fwd" : "tc_conv2_dgrad");
}

// ring depths: C = 128: 6 stages forward (+ 48 KB of staging), 9 data gradient; C = 256 (four halo chunks): 4 and 3
constexpr int kStF128 = 6, kStD128 = 9, kStF256 = 4, kStD256 = 3;

size_t conv2_smem(int C, bool fwd, int ksize, int dil) {
    const int half = ksize / 2, rows = C == 128 ? 256 : 128, chunks = C / 64;
    const size_t a0 = ((size_t)(rows + 2 * half) * 128 + 1023) / 1024 * 1024, a1 = ((size_t)(rows + 2 * half * dil) * 128 + 1023) / 1024 * 1024;
    const size_t am = a0 > a1 ? a0 : a1;
    const int st = C == 128 ? (fwd ? kStF128 : kStD128) : (fwd ? kStF256 : kStD256);
    return (size_t)((!fwd && C == 256) ? 2 : 1) * chunks * am + (size_t)st * kBTile + (fwd ? 8 * 3 * 2048 : 0) + 1024 + 256;
}

}  // namespace

// halo kernel covers: C = 128 (L a multiple of 256) or C = 256 (L a multiple of 128), odd kernel sizes whose halos fit shared memory
bool pb_tc_conv2_ok(int B, int L, int C, int ksize, int dil) {
    if (!((C == 128 || C == 256) && (ksize & 1) && ksize >= 3 && dil >= 1 && B >= 1)) return false;
    const int rows = C == 128 ? 256 : 128;
    if (L % rows != 0) return false;
    const int half = ksize / 2;
    if ((2 * half) % 8 != 0 || (2 * half * dil) % 8 != 0) return false;     // halo tails are whole 8-row swizzle groups
    return conv2_smem(C, true, ksize, dil) <= 232448 && conv2_smem(C, false, ksize, dil) <= 232448 && 2 * half * dil <= rows;
}

int pb_tc_conv2_fwd(const void* x, const void* wn, const void* ww, const float* bn, const float* bw, const float* g2l, void* zn,
                    void* zw, void* u, double* stats, int B, int L, int C, int ksize, int dil, int bo_mode, cudaStream_t st) {
    Conv2Params p{};
    p.x = (const bf16*)x; p.bn = bn; p.bw = bw; p.g2l = g2l;
    p.zn = (bf16*)zn; p.zw = (bf16*)zw; p.u = (bf16*)u; p.stats = stats;
    p.L = L; p.ksize = ksize; p.dil = dil; p.base_offset_mode = bo_mode;
    if (C == 256) return launch2<true, kStF256, 256>(x, x, wn, ww, p, B, L, st);
    return launch2<true, kStF128, 128>(x, x, wn, ww, p, B, L, st);
}

int pb_tc_conv2_dgrad(const void* du, const void* dzn, const void* dzw, const void* wn_bwd, const void* ww_bwd, void* dx, int B,
                      int L, int C, int ksize, int dil, int bo_mode, cudaStream_t st) {
    Conv2Params p{};
    p.du = (const bf16*)du; p.dx = (bf16*)dx;
    p.L = L; p.ksize = ksize; p.dil = dil; p.base_offset_mode = bo_mode;
    if (C == 256) return launch2<false, kStD256, 256>(dzn, dzw, wn_bwd, ww_bwd, p, B, L, st);
    return launch2<false, kStD128, 128>(dzn, dzw, wn_bwd, ww_bwd, p, B, L, st);
}
