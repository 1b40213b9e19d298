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

// one CTA per sample
__global__ void corrupt_pad_kernel(const uint8_t* __restrict__ seq_bytes, const int64_t* __restrict__ seq_off,
                                   const int64_t* __restrict__ crop_start, const uint8_t* __restrict__ tok_mask,
                                   const uint8_t* __restrict__ tok_rand, const int64_t* __restrict__ draw_off,
                                   const float* __restrict__ coin, const uint8_t* __restrict__ keep, const uint8_t* __restrict__ add,
                                   const uint8_t* __restrict__ ann, int64_t* __restrict__ x_local, int64_t* __restrict__ y_local,
                                   double* __restrict__ w_local, float* __restrict__ x_global, float* __restrict__ y_global,
                                   double* __restrict__ w_global, int L, int A) {
    __shared__ int any_ann;
    const int b = blockIdx.x;
    if (threadIdx.x == 0) any_ann = 0;
    __syncthreads();
    // ---- sequence: tokenise, <sos>/<eos>, crop, randomise, right-pad (:50-54,:78-83,:100-105,:165-167,:175)
    const int64_t s0 = seq_off[b];
    const int64_t len = seq_off[b + 1] - s0;      // residues
    const int64_t n_ids = len + 2;
    int64_t start = 0, n = n_ids;
    if (n_ids > L) { start = crop_start[b]; n = L; }
    const int64_t d0 = draw_off[b];
    for (int j = threadIdx.x; j < L; j += blockDim.x) {
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
    int local_any = 0;
    for (int a = threadIdx.x; a < A; a += blockDim.x) {
        const size_t o = (size_t)b * A + a;
        const int v = ann[o];
        local_any |= (v != 0);
        y_global[o] = (float)v;
        x_global[o] = noised ? (float)((v + (int)add[o]) * (int)keep[o]) : 0.f;
    }
    if (local_any) atomicOr(&any_ann, 1);
    __syncthreads();
    const double w = any_ann ? 1.0 : 0.0;
    for (int a = threadIdx.x; a < A; a += blockDim.x) w_global[(size_t)b * A + a] = w;
}

}  // namespace

extern "C" int pb_corrupt_pad(const uint8_t* seq_bytes, const int64_t* seq_off, const int64_t* crop_start,
                              const uint8_t* tok_mask, const uint8_t* tok_rand, const int64_t* draw_off, const float* coin,
                              const uint8_t* keep, const uint8_t* add, const uint8_t* ann, int64_t* x_local, int64_t* y_local,
                              double* w_local, float* x_global, float* y_global, double* w_global, int B, int L, int A,
                              void* stream) {
    PB_REQUIRE(B > 0 && L >= 2 && A > 0, PB_ERR_SHAPE, "corrupt_pad: bad shape B=%d L=%d A=%d", B, L, A);
    corrupt_pad_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(seq_bytes, seq_off, crop_start, tok_mask, tok_rand, draw_off, coin,
                                                            keep, add, ann, x_local, y_local, w_local, x_global, y_global,
                                                            w_global, L, A);
    pb_count_launch(1);
    return pb_check_launch("pb_corrupt_pad");
}
