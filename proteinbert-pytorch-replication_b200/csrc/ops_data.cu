// Input corruption and padding as one integer kernel (data_processing.py:30-183,337-348).
// Everything random is an INPUT (the draws the reference makes, in its order), so given the same draws the
// six output tensors are bit-identical to UniRefGO_PretrainingDataset.__getitem__ + default_collate.
#include "common.cuh"
#include "../../include/pb_b200.h"

namespace {

// create_amino_acid_vocab (data_processing.py:337-348): <pad>=0 <sos>=1 <eos>=2 <unk>=3, then the 22 letters
__device__ __forceinline__ int64_t residue_id(uint8_t ch) {
    const char* alphabet = "ACDEFGHIKLMNPQRSTUVWXY";
#pragma unroll
    for (int i = 0; i < 22; ++i)
        if ((uint8_t)alphabet[i] == ch) return 4 + i;
    return 3;   // set_default_index(<unk>)
}

// grid = (samples, slices): every CTA of a sample first folds the sample's whole annotation row into "any annotation set"
// (8.9 KB, read again by each of the slices: cheaper than a second launch or a grid-wide flag), then writes its slice of
// the sequence outputs and of the annotation outputs.  One CTA per sample (round 1) ran at 4.5 % of the HBM rate: 64 CTAs
// cannot pull 12 MB through 64 SMs' worth of load/store slots in the 2 us the traffic costs.
constexpr int kSlices = 16;

__global__ void __launch_bounds__(256) corrupt_pad_kernel(
    const uint8_t* __restrict__ seq_bytes, const int64_t* __restrict__ seq_off, const int64_t* __restrict__ crop_start,
    const uint8_t* __restrict__ tok_mask, const uint8_t* __restrict__ tok_rand, const int64_t* __restrict__ draw_off,
    const float* __restrict__ coin, const uint8_t* __restrict__ keep, const uint8_t* __restrict__ add,
    const uint8_t* __restrict__ ann, int64_t* __restrict__ x_local, int64_t* __restrict__ y_local, double* __restrict__ w_local,
    float* __restrict__ x_global, float* __restrict__ y_global, double* __restrict__ w_global, int L, int A) {
    __shared__ int any_ann;
    const int b = blockIdx.x, sl = blockIdx.y;
    if (threadIdx.x == 0) any_ann = 0;
    __syncthreads();
    {
        int local_any = 0;
        const uint8_t* row = ann + (size_t)b * A;
        for (int a = threadIdx.x; a < A; a += blockDim.x) local_any |= row[a];
        if (local_any) atomicOr(&any_ann, 1);
    }
    // ---- sequence: tokenise, <sos>/<eos>, crop, randomise, right-pad (:50-54,:78-83,:100-105,:165-167,:175)
    const int64_t s0 = seq_off[b];
    const int64_t len = seq_off[b + 1] - s0;      // residues
    const int64_t n_ids = len + 2;
    int64_t start = 0, n = n_ids;
    if (n_ids > L) { start = crop_start[b]; n = L; }
    const int64_t d0 = draw_off[b];
    const int l_per = (L + kSlices - 1) / kSlices;
    for (int j = sl * l_per + threadIdx.x; j < min(L, (sl + 1) * l_per); j += blockDim.x) {
        int64_t id = 0, xm = 0;
        if (j < n) {
            const int64_t k = start + j;            // index into [<sos>, residues..., <eos>]
            id = (k == 0) ? 1 : (k == n_ids - 1 ? 2 : residue_id(seq_bytes[s0 + k - 1]));
            const bool m = tok_mask[d0 + j] != 0 && id != 0 && id != 1 && id != 2;
            xm = m ? (int64_t)tok_rand[d0 + j] : id;
        }
        const size_t o = (size_t)b * L + j;
        y_local[o] = id;
        x_local[o] = xm;
        w_local[o] = (id != 0) ? 1.0 : 0.0;
    }
    // ---- annotations: (ann + add) * keep when the coin is > 0.5, zeros otherwise (:125-142,:173,:176)
    const bool noised = coin[b] > 0.5f;
    __syncthreads();
    const double w = any_ann ? 1.0 : 0.0;
    const int a_per = (A + kSlices - 1) / kSlices;
    for (int a = sl * a_per + threadIdx.x; a < min(A, (sl + 1) * a_per); a += blockDim.x) {
        const size_t o = (size_t)b * A + a;
        const int v = ann[o];
        y_global[o] = (float)v;
        x_global[o] = noised ? (float)((v + (int)add[o]) * (int)keep[o]) : 0.f;
        w_global[o] = w;
    }
}

// ---- f-3: the random draws of UniRefGO_PretrainingDataset.__getitem__ made on the device -------------------------------------
// Philox4x32-10 (Salmon et al., SC'11), counter-based: counter = (element index, stream, batch counter, sample), key = seed.
// The DISTRIBUTIONS are the reference's (data_processing.py:78-83,100-105,125-142); the stream is not torch's CPU generator,
// so this mode is statistically equivalent, not bit-identical, to the host-draw mode.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }      // [0, 1)

__global__ void __launch_bounds__(256) draw_batch_kernel(unsigned long long seed, unsigned long long counter,
                                                         const int64_t* __restrict__ seq_off, const int64_t* __restrict__ draw_off,
                                                         int64_t* __restrict__ crop_start, uint8_t* __restrict__ tok_mask,
                                                         uint8_t* __restrict__ tok_rand, float* __restrict__ coin,
                                                         uint8_t* __restrict__ keep, uint8_t* __restrict__ add, int L, int A, float p_tok,
                                                         float p_keep, float p_add) {
    const int b = blockIdx.x, sl = blockIdx.y;
    const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
    const uint32_t cnt = (uint32_t)counter;
    if (sl == 0 && threadIdx.x == 0) {                       // per-sample scalars: crop start (:78-83) and the coin (:130)
        const uint4 r = philox4x32_10(make_uint4(0u, 0u, cnt, (uint32_t)b), key);
        const int64_t n_ids = seq_off[b + 1] - seq_off[b] + 2;
        const int64_t span = n_ids - L;                      // torch.randint(0, len - max_length): [0, span)
        crop_start[b] = span > 0 ? (int64_t)(((unsigned long long)r.x * (unsigned long long)span) >> 32) : 0;
        coin[b] = u01(r.y);
    }
    // tokens: one Philox block -> two positions (mask draw, replacement id in [3, 26))
    const int64_t d0 = draw_off[b];
    const int n = (int)(draw_off[b + 1] - d0);
    const int pairs = (n + 1) / 2, pp = (pairs + kSlices - 1) / kSlices;
    for (int q = sl * pp + threadIdx.x; q < min(pairs, (sl + 1) * pp); q += blockDim.x) {
        const uint4 r = philox4x32_10(make_uint4((uint32_t)q, 1u, cnt, (uint32_t)b), key);
        const int j = 2 * q;
        tok_mask[d0 + j] = u01(r.x) < p_tok;
        tok_rand[d0 + j] = (uint8_t)(3u + (uint32_t)(((unsigned long long)r.y * 23ull) >> 32));
        if (j + 1 < n) {
            tok_mask[d0 + j + 1] = u01(r.z) < p_tok;
            tok_rand[d0 + j + 1] = (uint8_t)(3u + (uint32_t)(((unsigned long long)r.w * 23ull) >> 32));
        }
    }
    // annotations: one Philox block -> two annotations (keep w.p. p_keep, add w.p. p_add)
    const int apairs = (A + 1) / 2, ap = (apairs + kSlices - 1) / kSlices;
    for (int q = sl * ap + threadIdx.x; q < min(apairs, (sl + 1) * ap); q += blockDim.x) {
        const uint4 r = philox4x32_10(make_uint4((uint32_t)q, 2u, cnt, (uint32_t)b), key);
        const size_t o = (size_t)b * A + 2 * q;
        keep[o] = u01(r.x) < p_keep;
        add[o] = u01(r.y) < p_add;
        if (2 * q + 1 < A) {
            keep[o + 1] = u01(r.z) < p_keep;
            add[o + 1] = u01(r.w) < p_add;
        }
    }
}

}  // namespace

extern "C" int pb_corrupt_pad(const uint8_t* seq_bytes, const int64_t* seq_off, const int64_t* crop_start,
                              const uint8_t* tok_mask, const uint8_t* tok_rand, const int64_t* draw_off, const float* coin,
                              const uint8_t* keep, const uint8_t* add, const uint8_t* ann, int64_t* x_local, int64_t* y_local,
                              double* w_local, float* x_global, float* y_global, double* w_global, int B, int L, int A,
                              void* stream) {
    PB_REQUIRE(B > 0 && L >= 2 && A > 0, PB_ERR_SHAPE, "corrupt_pad: bad shape B=%d L=%d A=%d", B, L, A);
    corrupt_pad_kernel<<<dim3(B, kSlices), 256, 0, (cudaStream_t)stream>>>(seq_bytes, seq_off, crop_start, tok_mask, tok_rand, draw_off,
                                                                          coin, keep, add, ann, x_local, y_local, w_local, x_global,
                                                                          y_global, w_global, L, A);
    pb_count_launch(1);
    return pb_check_launch("pb_corrupt_pad");
}

extern "C" int pb_draw_batch(unsigned long long seed, unsigned long long counter, const int64_t* seq_off, const int64_t* draw_off,
                             int64_t* crop_start, uint8_t* tok_mask, uint8_t* tok_rand, float* coin, uint8_t* keep, uint8_t* add,
                             int B, int L, int A, float p_tok, float p_keep, float p_add, void* stream) {
    PB_REQUIRE(B > 0 && L >= 2 && A > 0, PB_ERR_SHAPE, "draw_batch: bad shape B=%d L=%d A=%d", B, L, A);
    PB_REQUIRE(seq_off && draw_off && crop_start && tok_mask && tok_rand && coin && keep && add, PB_ERR_SHAPE, "draw_batch: null argument");
    draw_batch_kernel<<<dim3(B, kSlices), 256, 0, (cudaStream_t)stream>>>(seed, counter, seq_off, draw_off, crop_start, tok_mask, tok_rand,
                                                                         coin, keep, add, L, A, p_tok, p_keep, p_add);
    pb_count_launch(1);
    return pb_check_launch("pb_draw_batch");
}
