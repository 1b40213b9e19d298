// Global-track operators (all float, M = batch rows): dense layers with split-K, and the
// residual + LayerNorm over G.  Reference call sites: modules.py:166-199,208,221-229,255-262,286-293.
#include <atomic>
#include "gemm_simt.cuh"
#include "../../include/pb_b200.h"

namespace {

constexpr int kMaxSplit = 32;

struct EpLinearDirect {
    const float* bias; float* Z; float* Y; int N; int act;
    __device__ void operator()(int m, int n, float acc, int, int) const {
        const float z = acc + (bias ? bias[n] : 0.f);
        const size_t i = (size_t)m * N + n;
        if (Z) Z[i] = z;
        Y[i] = act_fwd(act, z);
    }
};
struct EpPartial {
    float* ws; int M; int N;
    __device__ void operator()(int m, int n, float acc, int split, int) const {
        ws[((size_t)split * M + m) * N + n] = acc;
    }
};
struct EpStore {
    float* out; long ld;
    __device__ void operator()(int m, int n, float acc, int, int) const { out[(size_t)m * ld + n] = acc; }
};
struct EpStoreAcc {
    float* out; long ld; int accumulate;
    __device__ void operator()(int m, int n, float acc, int, int) const {
        const size_t i = (size_t)m * ld + n;
        out[i] = accumulate ? out[i] + acc : acc;
    }
};

__global__ void linear_reduce_kernel(const float* __restrict__ ws, const float* __restrict__ bias, float* Z, float* Y,
                                     int M, int N, int nsplit, int act) {
    const long long total = (long long)M * N;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int k = 0; k < nsplit; ++k) s += ws[(size_t)k * total + i];
        const float z = s + (bias ? bias[i % N] : 0.f);
        if (Z) Z[i] = z;
        Y[i] = act_fwd(act, z);
    }
}

__global__ void partial_reduce_kernel(const float* __restrict__ ws, float* out, long long total, int nsplit, int accumulate) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int k = 0; k < nsplit; ++k) s += ws[(size_t)k * total + i];
        out[i] = accumulate ? out[i] + s : s;
    }
}

// dZ = dY * act'(ZY); db[n] = sum_m dZ[m,n].  block = (32 columns, 8 row groups): the rows of a column are
// split over threadIdx.y and combined through shared memory, so the dependent-load chain is M/8 long.
__global__ void linear_dz_kernel(const float* __restrict__ dY, const float* __restrict__ ZY, float* __restrict__ dZ,
                                 float* __restrict__ db, int M, int N, int act) {
    __shared__ float part[8][33];
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

int pick_split(int tiles, int K) {
    int s = 296 / (tiles > 0 ? tiles : 1);
    if (s > kMaxSplit) s = kMaxSplit;
    while (s > 1 && K / s < 64) --s;
    return s < 1 ? 1 : s;
}

// y = LN(a + b + coef*c): one CTA per row, dynamic smem of G floats.
__global__ void res_ln_rows_fwd_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ c,
                                       const float* __restrict__ w_param, int Kdim, const float* __restrict__ gamma,
                                       const float* __restrict__ beta, float* __restrict__ xhat, float* __restrict__ rstd_out,
                                       float* __restrict__ y, int G) {
    extern __shared__ float row[];
    __shared__ float red[32];
    const int r = blockIdx.x;
    float coef = 0.f;
    if (c) {
        float t = 0.f;
        for (int k = threadIdx.x; k < Kdim; k += blockDim.x) t += w_param[k];
        coef = block_sum(t, red) / (float)Kdim;
    }
    float s = 0.f;
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        const size_t i = (size_t)r * G + g;
        float v = a[i];
        if (b) v += b[i];
        if (c) v += coef * c[i];
        row[g] = v;
        s += v;
    }
    const float mean = block_sum(s, red) / (float)G;
    float q = 0.f;
    for (int g = threadIdx.x; g < G; g += blockDim.x) { const float d = row[g] - mean; q += d * d; }
    const float var = block_sum(q, red) / (float)G;
    const float rstd = rsqrtf(var + 1e-5f);
    if (threadIdx.x == 0) rstd_out[r] = rstd;
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        const size_t i = (size_t)r * G + g;
        const float xh = (row[g] - mean) * rstd;
        xhat[i] = xh;
        y[i] = xh * gamma[g] + beta[g];
    }
}

__global__ void res_ln_rows_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ gamma,
                                       const float* __restrict__ xhat, const float* __restrict__ rstd,
                                       const float* __restrict__ c, const float* __restrict__ w_param, int Kdim,
                                       float* __restrict__ dx, float* __restrict__ dc, float* __restrict__ dgamma,
                                       float* __restrict__ dbeta, float* __restrict__ dw_param, int G) {
    __shared__ float red[32];
    const int r = blockIdx.x;
    float s1 = 0.f, s2 = 0.f;
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        const size_t i = (size_t)r * G + g;
        const float d = dy[i] * gamma[g];
        s1 += d;
        s2 += d * xhat[i];
    }
    const float m1 = block_sum(s1, red) / (float)G;
    const float m2 = block_sum(s2, red) / (float)G;
    float coef = 0.f;
    if (c) {
        float t = 0.f;
        for (int k = threadIdx.x; k < Kdim; k += blockDim.x) t += w_param[k];
        coef = block_sum(t, red) / (float)Kdim;
    }
    const float rs = rstd[r];
    float dcoef = 0.f;
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        const size_t i = (size_t)r * G + g;
        const float xh = xhat[i], dyv = dy[i];
        const float v = rs * (dyv * gamma[g] - m1 - xh * m2);
        dx[i] = v;
        if (c) {
            if (dc) dc[i] = coef * v;
            dcoef += v * c[i];
        }
        atomicAdd(&dgamma[g], dyv * xh);
        atomicAdd(&dbeta[g], dyv);
    }
    if (c && dw_param) {
        const float tot = block_sum(dcoef, red) / (float)Kdim;
        for (int k = threadIdx.x; k < Kdim; k += blockDim.x) atomicAdd(&dw_param[k], tot);
    }
}

}  // namespace

extern "C" {

int pb_linear_fwd(const float* X, const float* W, const float* bias, float* Z, float* Y, int M, int N, int K, int act,
                  void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(M > 0 && N > 0 && K > 0, PB_ERR_SHAPE, "linear_fwd: bad shape M=%d N=%d K=%d", M, N, K);
    cudaStream_t st = (cudaStream_t)stream;
    const int tiles = ceil_div(N, 64) * ceil_div(M, 64);
    int nsplit = pick_split(tiles, K);
    if (nsplit > 1 && (ws == nullptr || ws_bytes < (size_t)nsplit * M * N * sizeof(float))) nsplit = 1;
    RowMajorA<float> al{X, K};
    WeightNK<float> bl{W, K};
    if (nsplit == 1) {
        launch_gemm_simt<64, 64, 16, 4, 4>(al, bl, EpLinearDirect{bias, Z, Y, N, act}, M, N, K, 1, 1, st);
        pb_count_launch(1);
    } else {
        launch_gemm_simt<64, 64, 16, 4, 4>(al, bl, EpPartial{(float*)ws, M, N}, M, N, K, nsplit, 1, st);
        linear_reduce_kernel<<<min(ceil_div((long)M * N, 256), 1184), 256, 0, st>>>((const float*)ws, bias, Z, Y, M, N, nsplit, act);
        pb_count_launch(2);
    }
    return pb_check_launch("pb_linear_fwd");
}

int pb_linear_bwd(const float* dY, const float* ZY, const float* X, const float* W, float* dZ_buf, float* dX, float* dW,
                  float* db, int M, int N, int K, int act, int accumulate_dx, void* ws, size_t ws_bytes, void* stream) {
    PB_REQUIRE(M > 0 && N > 0 && K > 0, PB_ERR_SHAPE, "linear_bwd: bad shape M=%d N=%d K=%d", M, N, K);
    PB_REQUIRE(dZ_buf != nullptr, PB_ERR_WORKSPACE, "linear_bwd: dZ_buf is required");
    PB_REQUIRE(act == PB_ACT_NONE || ZY != nullptr, PB_ERR_SHAPE, "linear_bwd: ZY required for act=%d", act);
    cudaStream_t st = (cudaStream_t)stream;
    linear_dz_kernel<<<ceil_div(N, 32), dim3(32, 8), 0, st>>>(dY, ZY, dZ_buf, db, M, N, act);
    pb_count_launch(1);
    if (dX) {
        // dX[m,k] = sum_n dZ[m,n] W[n,k]
        const int tiles = ceil_div(K, 64) * ceil_div(M, 64);
        int nsplit = pick_split(tiles, N);
        if (nsplit > 1 && (ws == nullptr || ws_bytes < (size_t)nsplit * M * K * sizeof(float))) nsplit = 1;
        RowMajorA<float> al{dZ_buf, N};
        MatKN<float> bl{W, K};
        if (nsplit == 1) {
            launch_gemm_simt<64, 64, 16, 4, 4>(al, bl, EpStoreAcc{dX, K, accumulate_dx}, M, K, N, 1, 1, st);
            pb_count_launch(1);
        } else {
            launch_gemm_simt<64, 64, 16, 4, 4>(al, bl, EpPartial{(float*)ws, M, K}, M, K, N, nsplit, 1, st);
            partial_reduce_kernel<<<min(ceil_div((long)M * K, 256), 1184), 256, 0, st>>>((const float*)ws, dX, (long long)M * K,
                                                                                       nsplit, accumulate_dx);
            pb_count_launch(2);
        }
    }
    if (dW) {
        // dW[n,k] = sum_m dZ[m,n] X[m,k]
        ColMajorA<float> al{dZ_buf, N};
        MatKN<float> bl{X, K};
        launch_gemm_simt<64, 64, 16, 4, 4>(al, bl, EpStore{dW, K}, N, K, M, 1, 1, st);
        pb_count_launch(1);
    }
    return pb_check_launch("pb_linear_bwd");
}

// out[n, k] = sum_m A[m, n] * B[m, k]: the weight gradient dW = dZ^T X of a dense layer as a stand-alone product, for
// callers that take it off the critical path (the fused global-track backward kernels can skip their dW part).
int pb_matmul_tn(const float* A, const float* B, float* out, int M, int N, int K, void* stream) {
    PB_REQUIRE(M > 0 && N > 0 && K > 0 && A && B && out, PB_ERR_SHAPE, "matmul_tn: bad arguments M=%d N=%d K=%d", M, N, K);
    launch_gemm_simt<64, 64, 16, 4, 4>(ColMajorA<float>{A, N}, MatKN<float>{B, K}, EpStore{out, K}, N, K, M, 1, 1, (cudaStream_t)stream);
    pb_count_launch(1);
    return pb_check_launch("pb_matmul_tn");
}

int pb_res_ln_rows_fwd(const float* a, const float* b, const float* c, const float* w_param, int Kdim, const float* gamma,
                       const float* beta, float* xhat, float* rstd, float* y, int B, int G, void* stream) {
    PB_REQUIRE(B > 0 && G > 0 && G <= 12288, PB_ERR_SHAPE, "res_ln_rows_fwd: bad shape B=%d G=%d", B, G);
    PB_REQUIRE(c == nullptr || (w_param != nullptr && Kdim > 0), PB_ERR_SHAPE, "res_ln_rows_fwd: c needs w_param");
    res_ln_rows_fwd_kernel<<<B, 128, G * sizeof(float), (cudaStream_t)stream>>>(a, b, c, w_param, Kdim, gamma, beta, xhat,
                                                                               rstd, y, G);
    pb_count_launch(1);
    return pb_check_launch("pb_res_ln_rows_fwd");
}

int pb_res_ln_rows_bwd(const float* dy, const float* gamma, const float* xhat, const float* rstd, const float* c,
                       const float* w_param, int Kdim, float* dx, float* dc, float* dgamma, float* dbeta, float* dw_param,
                       int B, int G, void* stream) {
    PB_REQUIRE(B > 0 && G > 0, PB_ERR_SHAPE, "res_ln_rows_bwd: bad shape B=%d G=%d", B, G);
    cudaStream_t st = (cudaStream_t)stream;
    PB_CUDA(cudaMemsetAsync(dgamma, 0, G * sizeof(float), st));
    PB_CUDA(cudaMemsetAsync(dbeta, 0, G * sizeof(float), st));
    if (c && dw_param) PB_CUDA(cudaMemsetAsync(dw_param, 0, Kdim * sizeof(float), st));
    res_ln_rows_bwd_kernel<<<B, 128, 0, st>>>(dy, gamma, xhat, rstd, c, w_param, Kdim, dx, dc, dgamma, dbeta, dw_param, G);
    pb_count_launch(1);
    return pb_check_launch("pb_res_ln_rows_bwd");
}

}  // extern "C"

// =================================================================================================
// pb_linear_fwd_noalloc: pb_linear_fwd for callers that have no scratch buffer to hand over.
//
// Who needs this: host programs that must not (or cannot) create device allocations at the point where
// they call into this library -- for example an application that gave the whole device heap to its own
// memory arena at start-up, a serving loop that cannot afford the device-wide synchronisation a
// cudaMalloc/cudaFree implies, or a plugin host that lets plugins launch kernels but not allocate.
// pb_linear_fwd needs split-K scratch from its caller; this variant keeps that scratch inside the
// library image instead, in the fixed-size device array below, so the call itself allocates nothing
// (no cudaMalloc/cudaFree, no allocating library routine).  Inputs and outputs are still the caller's.
//
// Cost and limits: every GPU that loads libpb_b200.so holds the 4 MiB array until the process exits.
// One call may use at most PB_NOALLOC_MAX_MN = 32768 output elements (M*N; e.g. batch 64 x 512 features,
// the annotation->global layer of the default model); larger problems are refused with PB_ERR_UNSUPPORTED.
//
// Sharing: all calls on a device use the same array.  Nothing in it has to start at zero -- every partial
// sum that the reduction reads is overwritten by the GEMM kernel of the same call first -- so stale
// contents from an earlier call are harmless; but two calls that run concurrently on one GPU (different
// streams, or different host threads) would overwrite each other's partial sums: serialise them.
//
// Set-up: the array's device address is looked up once per device with cudaGetSymbolAddress, by the first
// call made while that device is current (the array's name is only a host-side handle, it is not usable as
// a pointer in a launch).  Make that first call before starting any stream capture on that device: with lazy module
// loading the look-up may load this library's device code onto the GPU, which is not permitted during capture.  The
// address is cached per device index of the CURRENT device, so `stream` must belong to the current device, and the
// cache does not survive a cudaDeviceReset.
// =================================================================================================
#define PB_NOALLOC_MAX_MN 32768
#define PB_NOALLOC_SCRATCH_FLOATS (kMaxSplit * PB_NOALLOC_MAX_MN)          /* 32 splits x 32768 floats = 4 MiB */
#define PB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) float g_linear_scratch[PB_NOALLOC_SCRATCH_FLOATS];
static std::atomic<float*> g_linear_scratch_addr[PB_NOALLOC_MAX_DEVICES];      // zero-initialised; filled once per device

extern "C" int pb_linear_fwd_noalloc(const float* X, const float* W, const float* bias, float* Z, float* Y, int M, int N, int K,
                                     int act, void* stream) {
    PB_REQUIRE(M > 0 && N > 0 && K > 0, PB_ERR_SHAPE, "linear_fwd_noalloc: bad shape M=%d N=%d K=%d", M, N, K);
    PB_REQUIRE((long long)M * N <= PB_NOALLOC_MAX_MN, PB_ERR_UNSUPPORTED,
               "linear_fwd_noalloc: M*N = %lld exceeds the built-in scratch limit of %d output elements; "
               "use pb_linear_fwd with a caller-provided workspace", (long long)M * N, PB_NOALLOC_MAX_MN);
    int dev = -1;
    PB_CUDA(cudaGetDevice(&dev));
    PB_REQUIRE(dev >= 0 && dev < PB_NOALLOC_MAX_DEVICES, PB_ERR_UNSUPPORTED, "linear_fwd_noalloc: device index %d not supported", dev);
    float* ws = g_linear_scratch_addr[dev].load(std::memory_order_acquire);
    if (ws == nullptr) {                                    // once per device, with that device current
        float* ptr = nullptr;
        PB_CUDA(cudaGetSymbolAddress((void**)&ptr, g_linear_scratch));
        PB_REQUIRE(ptr != nullptr, PB_ERR_CUDA, "linear_fwd_noalloc: cudaGetSymbolAddress returned a null address");
        g_linear_scratch_addr[dev].store(ptr, std::memory_order_release);   // racing threads store the same value
        ws = ptr;                                           // a real device address; never the symbol name itself
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int tiles = ceil_div(N, 64) * ceil_div(M, 64);
    const int nsplit = pick_split(tiles, K);                // <= kMaxSplit, so nsplit*M*N floats always fit
    RowMajorA<float> al{X, K};
    WeightNK<float> bl{W, K};
    if (nsplit == 1) {
        launch_gemm_simt<64, 64, 16, 4, 4>(al, bl, EpLinearDirect{bias, Z, Y, N, act}, M, N, K, 1, 1, st);
        pb_count_launch(1);
    } else {
        launch_gemm_simt<64, 64, 16, 4, 4>(al, bl, EpPartial{ws, M, N}, M, N, K, nsplit, 1, st);
        linear_reduce_kernel<<<min(ceil_div((long)M * N, 256), 1184), 256, 0, st>>>(ws, bias, Z, Y, M, N, nsplit, act);
        pb_count_launch(2);
    }
    return pb_check_launch("pb_linear_fwd_noalloc");
}
