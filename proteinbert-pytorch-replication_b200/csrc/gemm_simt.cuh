// FP32 SIMT GEMM with functor operand loaders and a functor epilogue.
//
// This is the exact-arithmetic engine: fp32 FMA with fp32 accumulation, used for
// the fp32 parity mode of every contraction on the path and, in both modes, for
// the small global-track GEMMs (M = batch).  The bf16 tensor-core kernels
// (tc_*.cu) replace it for the large local-track contractions.
//
//   C[m,n] = sum_k A(m,k) * B(k,n)        m<M, n<N, k in this CTA's K-chunk
//
// AL : float operator()(int m, int k)  + static constexpr bool kContig (k fastest in memory)
// BL : float operator()(int k, int n)  + static constexpr bool kContig
// EP : void  operator()(int m, int n, float acc, int split)
// gridDim = (ceil(N/BN), ceil(M/BM), splits * batch); blockIdx.z is handed to the
// loaders/epilogue through `zidx` so one launch can cover several taps/splits.
#pragma once
#include "common.cuh"

template <int BM, int BN, int BK, int TM, int TN, class AL, class BL, class EP>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_simt_kernel(AL al, BL bl, EP ep, int M, int N, int K, int kchunk, int nsplit) {
    constexpr int NT = (BM / TM) * (BN / TN);
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int tid = threadIdx.x;
    const int tx = tid % (BN / TN), ty = tid / (BN / TN);
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int zbatch = blockIdx.z / nsplit, split = blockIdx.z % nsplit;
    const int kbeg = split * kchunk;
    const int kend = min(K, kbeg + kchunk);

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    for (int k0 = kbeg; k0 < kend; k0 += BK) {
#pragma unroll
        for (int idx = tid; idx < BM * BK; idx += NT) {
            int m, k;
            if (AL::kContig) { m = idx / BK; k = idx % BK; } else { k = idx / BM; m = idx % BM; }
            const int gm = m0 + m, gk = k0 + k;
            As[k][m] = (gm < M && gk < kend) ? al(gm, gk, zbatch) : 0.f;
        }
#pragma unroll
        for (int idx = tid; idx < BN * BK; idx += NT) {
            int n, k;
            if (BL::kContig) { n = idx / BK; k = idx % BK; } else { k = idx / BN; n = idx % BN; }
            const int gn = n0 + n, gk = k0 + k;
            Bs[k][n] = (gn < N && gk < kend) ? bl(gk, gn, zbatch) : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; ++i) a[i] = As[kk][ty * TM + i];
#pragma unroll
            for (int j = 0; j < TN; ++j) b[j] = Bs[kk][tx * TN + j];
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int gm = m0 + ty * TM + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int gn = n0 + tx * TN + j;
            if (gn < N) ep(gm, gn, acc[i][j], split, zbatch);
        }
    }
}

template <int BM, int BN, int BK, int TM, int TN, class AL, class BL, class EP>
static inline void launch_gemm_simt(AL al, BL bl, EP ep, int M, int N, int K, int nsplit, int zbatch,
                                    cudaStream_t st) {
    if (M <= 0 || N <= 0) return;
    if (nsplit < 1) nsplit = 1;
    int kchunk = (K + nsplit - 1) / nsplit;
    kchunk = ((kchunk + BK - 1) / BK) * BK;
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, nsplit * zbatch);
    gemm_simt_kernel<BM, BN, BK, TM, TN, AL, BL, EP><<<grid, (BM / TM) * (BN / TN), 0, st>>>(
        al, bl, ep, M, N, K, kchunk, nsplit);
}

// ---- common loaders -----------------------------------------------------------
// row-major [rows, ld] read as A(m,k) = p[m*ld + k]
template <typename T>
struct RowMajorA {
    static constexpr bool kContig = true;
    const T* p; long ld;
    __device__ float operator()(int m, int k, int) const { return ldf(p, (size_t)m * ld + k); }
};
// row-major [K rows, ld] read transposed as A(m,k) = p[k*ld + m]
template <typename T>
struct ColMajorA {
    static constexpr bool kContig = false;
    const T* p; long ld;
    __device__ float operator()(int m, int k, int) const { return ldf(p, (size_t)k * ld + m); }
};
// weight stored [N, K] (nn.Linear layout): B(k,n) = p[n*ld + k]
template <typename T>
struct WeightNK {
    static constexpr bool kContig = true;
    const T* p; long ld;
    __device__ float operator()(int k, int n, int) const { return ldf(p, (size_t)n * ld + k); }
};
// matrix stored [K, N]: B(k,n) = p[k*ld + n]
template <typename T>
struct MatKN {
    static constexpr bool kContig = false;
    const T* p; long ld;
    __device__ float operator()(int k, int n, int) const { return ldf(p, (size_t)k * ld + n); }
};
