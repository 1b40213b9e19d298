// EXPERIMENT, not built into libpb_b200.so (round 2).  Result on B200 at B=64, L=512, C=128: 35.5 us (two GELU branches) and
// 24.5 us (one) against 6.8 + 17.4 us and 6.8 + 11.8 us for the two-launch form it was meant to replace: the LayerNorm-backward
// kernels are bound by instruction issue (~50 instructions per element, three quarters of them the two gelu' polynomials), not by
// the second read of dy and u that this kernel removes, and one 512-thread CTA per SM hides latency worse than two 256-thread
// CTAs.  Kept for the record; parity-tested (tests/test_gpu_parity.py::test_fused_layernorm_gelu_backward passed with it).
// Joint (L,C) LayerNorm backward + the GELU backwards behind it as ONE launch (modules.py:212,217 and the GELUs of :207,:215
// in reverse), bf16 activations.
//
// The two-launch form (ops_local.cu: ln_bwd_sums_vec_kernel, then ln_bwd_apply_gelu_kernel) reads dy and u twice: the
// input gradient of a LayerNorm over a whole sample needs two per-sample sums (sum dy*gamma, sum dy*gamma*xhat) over all L*C
// elements before the first output element can be formed, and a sample (128 KB of dy + 128 KB of u at L*C = 65536) does not
// fit one CTA together with the per-(l,c) accumulators of dgamma / dbeta.  Here a thread-block CLUSTER of 8 CTAs spans the
// L*C axis: CTA r owns a contiguous 1/8 of every sample, brings its slice of dy and u for a round of S samples into shared
// memory with bulk copies (one elected thread, no register staging), forms its partial sums, and the eight partials meet
// through distributed shared memory between two barrier.cluster -- then the same shared-memory copy of dy and u feeds the
// output pass.  dy and u cross HBM/L2 once; 12 launches and 2 x 16.8 MB of reads per block leave the step.
// A thread owns V (l,c) vectors of 8 channels and walks the cluster's samples, so dgamma/dbeta (summed over the batch) and
// the per-channel bias gradients stay in registers until the end, as in the two-launch form.
#include <stdlib.h>
#include "tc_common.cuh"
#include "tc_api.cuh"

namespace {

constexpr int kLC = 8;             // CTAs per cluster (portable size)
constexpr int kLT = 512;           // threads per CTA

struct LnBwdP {
    const bf16* dy; const bf16* u; const bf16* z0; const bf16* z1;
    const double* stats; const float* gamma;
    bf16* du; bf16* dz0; bf16* dz1;
    float* dgamma; float* dbeta; float* dsample; float* db0; float* db1;
    int B, C; long long n; float eps; int per_cluster;
};

__device__ __forceinline__ uint32_t lc_cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void lc_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ double lc_ld_remote_f64(const double* local, uint32_t rank) {
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local)), "r"(rank));
    double v;
    asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(v) : "r"(ra) : "memory");
    return v;
}
__device__ __forceinline__ void lc_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// per-thread asynchronous 16-byte copy global -> shared (the issuing thread alone reads the destination, after its wait_group)
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldf8g(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void unpack8(const uint4& r, float (&v)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
}
// stores v rounded to bf16 and leaves the rounded values in v
__device__ __forceinline__ void st8(bf16* p, float (&v)[8]) {
    uint4 r;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
    *reinterpret_cast<uint4*>(p) = r;
}

// shared memory: [S][V*512] uint4 dy | same for u | z ring [2 stages][1 or 2 tensors][V*512] uint4 | part[(2 + per_cluster)][C] floats | red[16][2S] doubles | cta_part[2S] doubles
// | tot[2S] doubles | prm[S][4] floats | mbarrier
template <bool TWO, int V>
__global__ void __launch_bounds__(kLT, 1) ln_bwd_cluster_kernel(const LnBwdP p) {
    constexpr int S = 8 / V;                     // samples per round: S * V * 512 vectors of dy and of u = 128 KB
    constexpr int kSlice = V * kLT;              // vectors of one sample owned by this CTA
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint4* sdy = reinterpret_cast<uint4*>(smem_raw);
    uint4* su = sdy + S * kSlice;
    constexpr int NZ = TWO ? 2 : 1;
    uint4* zring = su + S * kSlice;              // two samples of z0 (and z1) in flight per thread, copied by the thread that uses them
    float* part = reinterpret_cast<float*>(zring + 2 * NZ * kSlice);
    double* red = reinterpret_cast<double*>(part + (size_t)(2 + p.per_cluster) * p.C + ((2 + p.per_cluster) * p.C & 1));
    double* cta_part = red + 16 * 2 * S;
    double* tot = cta_part + 2 * S;
    float* prm = reinterpret_cast<float*>(tot + 2 * S);
    uint64_t* bar = reinterpret_cast<uint64_t*>(prm + 4 * S);

    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const uint32_t rank = lc_cluster_rank();
    const int cid = blockIdx.x / kLC;
    const int b0 = cid * p.per_cluster;
    const int nb = min(p.per_cluster, p.B - b0);          // samples of this cluster (may be <= 0 for a trailing cluster)
    const int C = p.C, vpr = C / 8, tv = lane % vpr;
    const long long n = p.n;
    const long long slice0 = (long long)rank * kSlice;    // first vector of this CTA's slice

    if (t == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    for (int e = t; e < (2 + p.per_cluster) * C; e += kLT) part[e] = 0.f;
    float g[V][8], ag[V][8], ab[V][8], c0[8], c1[8];
#pragma unroll
    for (int v = 0; v < V; ++v) {
        ldf8g(p.gamma + (slice0 + t + v * kLT) * 8, g[v]);          // parameters: not written by the previous kernel
#pragma unroll
        for (int j = 0; j < 8; ++j) { ag[v][j] = 0.f; ab[v][j] = 0.f; }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { c0[j] = 0.f; c1[j] = 0.f; }
    __syncthreads();
    pdl_wait();

    uint32_t parity = 0;
    for (int r0 = 0; r0 < nb; r0 += S) {
        const int ns = min(S, nb - r0);
        // ---- this round's slices of dy and u -> shared memory
        if (t == 0) {
            mbar_expect_tx(bar, (uint32_t)(ns * 2 * kSlice * 16));
            for (int s = 0; s < ns; ++s) {
                const size_t off = (size_t)(b0 + r0 + s) * n + (size_t)slice0 * 8;
                lc_bulk_g2s(sdy + s * kSlice, p.dy + off, kSlice * 16, bar);
                lc_bulk_g2s(su + s * kSlice, p.u + off, kSlice * 16, bar);
            }
        }
        // z of the round's first two samples: requested now, used after the sums (pass 2); one commit group per sample
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            if (s < ns) {
                const size_t off = (size_t)(b0 + r0 + s) * n + (size_t)(slice0 + t) * 8;
#pragma unroll
                for (int v = 0; v < V; ++v) {
                    cp_async16(zring + (s * NZ) * kSlice + v * kLT + t, p.z0 + off + (size_t)v * kLT * 8);
                    if (TWO) cp_async16(zring + (s * NZ + 1) * kSlice + v * kLT + t, p.z1 + off + (size_t)v * kLT * 8);
                }
            }
            cp_async_commit();
        }
        mbar_wait(bar, parity);
        parity ^= 1;
        // ---- pass 1: partial sums  s1 = sum dy*gamma, s2 = sum dy*gamma*u  per sample
        float s1[S], s2[S];
#pragma unroll
        for (int s = 0; s < S; ++s) {
            s1[s] = 0.f; s2[s] = 0.f;
            if (s < ns) {
#pragma unroll
                for (int v = 0; v < V; ++v) {
                    float d[8], x[8];
                    unpack8(sdy[s * kSlice + v * kLT + t], d);
                    unpack8(su[s * kSlice + v * kLT + t], x);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float a = d[j] * g[v][j];
                        s1[s] += a;
                        s2[s] = fmaf(a, x[j], s2[s]);
                    }
                }
            }
        }
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const double a = warp_sum_d((double)s1[s]), c = warp_sum_d((double)s2[s]);
            if (lane == 0) { red[warp * 2 * S + 2 * s] = a; red[warp * 2 * S + 2 * s + 1] = c; }
        }
        __syncthreads();
        if (t < 2 * S) {
            double a = 0.0;
#pragma unroll
            for (int w = 0; w < kLT / 32; ++w) a += red[w * 2 * S + t];
            cta_part[t] = a;
        }
        lc_cluster_sync();                                  // every CTA's partials are in its shared memory
        if (t < 2 * S) {
            double a = 0.0;
#pragma unroll
            for (int r = 0; r < kLC; ++r) a += lc_ld_remote_f64(cta_part + t, (uint32_t)r);
            tot[t] = a;
        }
        lc_cluster_sync();                                  // ... and have been read: cta_part may be rewritten next round
        if (t < ns) {
            const int b = b0 + r0 + t;
            const LnStat st = ln_stat(p.stats, b, (double)n, p.eps);
            const double mean = p.stats[2 * b] / (double)n;
            prm[4 * t] = st.mean; prm[4 * t + 1] = st.rstd;
            prm[4 * t + 2] = (float)(tot[2 * t] / (double)n);
            double var = p.stats[2 * b + 1] / (double)n - mean * mean;
            if (var < 0.0) var = 0.0;
            prm[4 * t + 3] = (float)((tot[2 * t + 1] - mean * tot[2 * t]) * rsqrt(var + (double)p.eps) / (double)n);
        }
        __syncthreads();
        // ---- pass 2: du, dz0, dz1 and the sums, dy and u from shared memory, z from global (next sample requested early)
        for (int s = 0; s < ns; ++s) {
            const float mean = prm[4 * s], rstd = prm[4 * s + 1], m1 = prm[4 * s + 2], m2 = prm[4 * s + 3];
            const size_t off = (size_t)(b0 + r0 + s) * n + (size_t)(slice0 + t) * 8;
            float osum[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) osum[j] = 0.f;
            cp_async_wait<1>();                   // this sample's z has landed (the next sample's may still be in flight)
            const uint4* zs = zring + ((s & 1) * NZ) * kSlice;
#pragma unroll
            for (int v = 0; v < V; ++v) {
                float d[8], x[8], za[8], zb[8], o[8], e[8];
                unpack8(sdy[s * kSlice + v * kLT + t], d);
                unpack8(su[s * kSlice + v * kLT + t], x);
                unpack8(zs[v * kLT + t], za);
                if (TWO) unpack8(zs[kSlice + v * kLT + t], zb);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float xh = (x[j] - mean) * rstd;
                    o[j] = rstd * (d[j] * g[v][j] - m1 - xh * m2);
                    ag[v][j] = fmaf(d[j], xh, ag[v][j]);
                    ab[v][j] += d[j];
                }
                st8(p.du + off + (size_t)v * kLT * 8, o);              // o now holds the stored (rounded) values
#pragma unroll
                for (int j = 0; j < 8; ++j) { e[j] = o[j] * gelu_grad_fast(za[j]); osum[j] += o[j]; }
                st8(p.dz0 + off + (size_t)v * kLT * 8, e);
#pragma unroll
                for (int j = 0; j < 8; ++j) c0[j] += e[j];
                if (TWO) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) e[j] = o[j] * gelu_grad_fast(zb[j]);
                    st8(p.dz1 + off + (size_t)v * kLT * 8, e);
#pragma unroll
                    for (int j = 0; j < 8; ++j) c1[j] += e[j];
                }
            }
            // the slots just read take the z of sample s + 2 (this thread's own slots: no other thread touches them)
            if (s + 2 < ns) {
                const size_t off2 = off + 2 * (size_t)n;
#pragma unroll
                for (int v = 0; v < V; ++v) {
                    cp_async16(zring + ((s & 1) * NZ) * kSlice + v * kLT + t, p.z0 + off2 + (size_t)v * kLT * 8);
                    if (TWO) cp_async16(zring + ((s & 1) * NZ + 1) * kSlice + v * kLT + t, p.z1 + off2 + (size_t)v * kLT * 8);
                }
            }
            cp_async_commit();
            if (p.dsample != nullptr) {          // dsample[b, c] = sum_l du: fold the rows this warp holds, then shared-memory adds
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    for (int sh = 16; sh >= vpr; sh >>= 1) osum[j] += __shfl_xor_sync(0xffffffffu, osum[j], sh);
                if (lane < vpr) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) atomicAdd(&part[(size_t)(2 + r0 + s) * C + tv * 8 + j], osum[j]);
                }
            }
        }
        cp_async_wait<0>();
        __syncthreads();          // all reads of this round's shared-memory tiles are done before the next round's copies land
    }
    if (nb > 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            for (int sh = 16; sh >= vpr; sh >>= 1) {
                c0[j] += __shfl_xor_sync(0xffffffffu, c0[j], sh);
                if (TWO) c1[j] += __shfl_xor_sync(0xffffffffu, c1[j], sh);
            }
        if (lane < vpr) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(&part[tv * 8 + j], c0[j]);
                if (TWO) atomicAdd(&part[C + tv * 8 + j], c1[j]);
            }
        }
#pragma unroll
        for (int v = 0; v < V; ++v) {
            const long long i = slice0 + t + v * kLT;
            atomicAdd(reinterpret_cast<float4*>(p.dgamma + i * 8), make_float4(ag[v][0], ag[v][1], ag[v][2], ag[v][3]));
            atomicAdd(reinterpret_cast<float4*>(p.dgamma + i * 8 + 4), make_float4(ag[v][4], ag[v][5], ag[v][6], ag[v][7]));
            atomicAdd(reinterpret_cast<float4*>(p.dbeta + i * 8), make_float4(ab[v][0], ab[v][1], ab[v][2], ab[v][3]));
            atomicAdd(reinterpret_cast<float4*>(p.dbeta + i * 8 + 4), make_float4(ab[v][4], ab[v][5], ab[v][6], ab[v][7]));
        }
        __syncthreads();
        const int last = p.dsample != nullptr ? 2 + nb : 2;
        for (int e = t; e < last * C; e += kLT) {
            const int slot = e / C, c = e - slot * C;
            if (slot == 1 && !TWO) continue;
            float* dst = slot == 0 ? p.db0 + c : slot == 1 ? p.db1 + c : p.dsample + (size_t)(b0 + slot - 2) * C + c;
            atomicAdd(dst, part[e]);
        }
    }
}

int g_state[2][2] = {{0, 0}, {0, 0}};        // [TWO][V == 2]: 0 unknown, 1 ok, -1 unavailable
int g_max_clusters[2][2] = {{0, 0}, {0, 0}};

size_t smem_bytes(int per_cluster, int C) {
    size_t part = (size_t)(2 + per_cluster) * C;
    part += part & 1;
    return (size_t)2 * 8 * kLT * 16 + (size_t)2 * 2 * 2 * kLT * 16 + part * 4 + (size_t)(16 * 16 + 16 + 16) * 8 + 8 * 4 * 4 + 16;
}

template <bool TWO, int V>
int prepare(size_t smem) {
    int& state = g_state[TWO][V == 2];
    if (state == 0) {
        auto kern = ln_bwd_cluster_kernel<TWO, V>;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) != cudaSuccess) {
            cudaGetLastError();
            state = -1;
            return state;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(kLC); cfg.blockDim = dim3(kLT); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = kLC; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        int nmax = 0;
        if (cudaOccupancyMaxActiveClusters(&nmax, kern, &cfg) != cudaSuccess) { cudaGetLastError(); nmax = 0; }
        g_max_clusters[TWO][V == 2] = nmax;
        state = nmax >= 1 ? 1 : -1;
    }
    return state;
}

template <bool TWO, int V>
int launch(const LnBwdP& p0, cudaStream_t st) {
    LnBwdP p = p0;
    constexpr int S = 8 / V;
    if (prepare<TWO, V>(smem_bytes(S, p.C)) != 1) return -1;
    const int nmax = g_max_clusters[TWO][V == 2];
    // as many clusters as fit at once (one wave), each walking per_cluster samples in rounds of S
    int per = (p.B + nmax - 1) / nmax;
    if (per < 1) per = 1;
    if (per < S && p.B >= S * nmax / 2) per = S;            // full rounds when the batch is large enough to still fill the GPU
    const int nclusters = (p.B + per - 1) / per;
    p.per_cluster = per;
    const size_t smem = smem_bytes(per, p.C);
    if (smem > 220 * 1024) return -1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kLC * nclusters); cfg.blockDim = dim3(kLT); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kLC; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pb_pdl_enabled() ? 2 : 1;
    PB_CUDA(cudaLaunchKernelEx(&cfg, ln_bwd_cluster_kernel<TWO, V>, p));
    pb_count_launch(1);
    return pb_check_launch("pb_ln_joint_bwd_gelu (cluster)");
}

int vectors_per_thread(int L, int C) {
    const long long nv = (long long)L * C / 8;
    if ((long long)L * C % 8) return 0;
    if (nv == (long long)kLC * kLT) return 1;
    if (nv == (long long)kLC * kLT * 2) return 2;
    return 0;
}

}  // namespace

// L*C = 32768 or 65536 elements per sample (the default model at sequence length 256 / 512), C/8 a divisor of 32
bool pb_ln_bwd_cluster_ok(int B, int L, int C) {
    static int enabled = -1;
    if (enabled < 0) {
        const char* e = getenv("PB_B200_LN_CLUSTER");
        enabled = (e == nullptr || atoi(e) != 0) ? 1 : 0;
    }
    if (!enabled || B <= 0 || C % 8 != 0 || C / 8 > 32 || 32 % (C / 8) != 0) return false;
    return vectors_per_thread(L, C) != 0;
}

// returns -1 when the cluster launch is unavailable on this device (the caller then takes the two-launch path)
int pb_ln_bwd_cluster(const void* dy, const void* u, const double* stats, const float* gamma, const void* z0, const void* z1, void* du,
                      void* dz0, void* dz1, float* dgamma, float* dbeta, float* dsample, float* db0, float* db1, int B, int L, int C,
                      float eps, cudaStream_t st) {
    LnBwdP p{(const bf16*)dy, (const bf16*)u, (const bf16*)z0, (const bf16*)z1, stats, gamma, (bf16*)du, (bf16*)dz0, (bf16*)dz1,
             dgamma, dbeta, dsample, db0, db1, B, C, (long long)L * C, eps, 0};
    const int V = vectors_per_thread(L, C);
    if (V == 1) return z1 ? launch<true, 1>(p, st) : launch<false, 1>(p, st);
    if (V == 2) return z1 ? launch<true, 2>(p, st) : launch<false, 2>(p, st);
    return -1;
}
