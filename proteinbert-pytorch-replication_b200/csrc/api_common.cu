// Library plumbing: error string, device check, launch counter, parameter staging kernels.
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <atomic>
#include "common.cuh"
#include "../../include/pb_b200.h"

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void pb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::atomic<int> g_pdl{-1};
bool pb_pdl_enabled() {
    int v = g_pdl.load(std::memory_order_relaxed);
    if (v < 0) {
        const char* e = getenv("PB_B200_PDL");
        v = (e == nullptr || e[0] != '0') ? 1 : 0;
        g_pdl.store(v);
    }
    return v != 0;
}

void pb_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int pb_check_launch(const char* what) {
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
        pb_set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
        (void)cudaGetLastError();
        return PB_ERR_CUDA;
    }
    return PB_OK;
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, long long n) {
    long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 3 < n) {
        const float4 v = *reinterpret_cast<const float4*>(src + i);
        __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&lo);
        o.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(dst + i) = o;
    } else {
        for (; i < n; ++i) dst[i] = __float2bfloat16_rn(src[i]);
    }
}

template <typename T>
__global__ void conv_pack_kernel(const float* __restrict__ w, T* __restrict__ fwd, T* __restrict__ bwd, int C, int k) {
    // w [co][ci][t] -> fwd [t][co][ci], bwd [t][ci][co]
    const long long n = (long long)C * C * k;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i % k);
        const int ci = (int)((i / k) % C);
        const int co = (int)(i / ((long long)k * C));
        const float v = w[i];
        if (fwd) stf(fwd, ((size_t)t * C + co) * C + ci, v);
        if (bwd) stf(bwd, ((size_t)t * C + ci) * C + co, v);
    }
}

__device__ __forceinline__ void st2(float* p, size_t i, float a, float b) { *reinterpret_cast<float2*>(p + i) = make_float2(a, b); }
__device__ __forceinline__ void st2(bf16* p, size_t i, float a, float b) { *reinterpret_cast<__nv_bfloat162*>(p + i) = __floats2bfloat162_rn(a, b); }

// One launch stages every weight of the local track: blockIdx.y walks a device table of jobs.
//   kind 0: src [d0, d1] row-major -> dst0[r*ld0 + c] (copy/cast) and dst1[c*ld1 + r] (transposed copy/cast)
//   kind 1: src [Co=d0, Ci=d1, k=d2] -> dst0 [k][Co][Ci], dst1 [k][Ci][Co]
template <typename T>
__global__ void stage_weights_kernel(const pb_stage_desc* __restrict__ jobs) {
    const pb_stage_desc j = jobs[blockIdx.y];
    T* d0 = (T*)j.dst0;
    T* d1 = (T*)j.dst1;
    if (j.kind == 0) {
        const long long n = (long long)j.d0 * j.d1;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
            const int r = (int)(i / j.d1), c = (int)(i % j.d1);
            const float v = j.src[i];
            if (d0) stf(d0, (size_t)r * j.ld0 + c, v);
            if (d1) stf(d1, (size_t)c * j.ld1 + r, v);
        }
    } else {
        const int C0 = j.d0, C1 = j.d1, k = j.d2;
        const long long n = (long long)C0 * C1 * k;
        if (((C0 | C1) & 1) == 0) {
            // gather form: threads walk the OUTPUT in pairs (coalesced 4/8-byte stores); the strided reads stay in L1/L2
            for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n / 2; i += (long long)gridDim.x * blockDim.x) {
                const long long o = 2 * i;
                if (d0) {
                    const int ci = (int)(o % C1), co = (int)((o / C1) % C0), t = (int)(o / ((long long)C1 * C0));
                    const float* sp = j.src + ((size_t)co * C1 + ci) * k + t;
                    st2(d0, (size_t)o, __ldg(sp), __ldg(sp + k));
                }
                if (d1) {
                    const int co = (int)(o % C0), ci = (int)((o / C0) % C1), t = (int)(o / ((long long)C1 * C0));
                    const float* sp = j.src + ((size_t)co * C1 + ci) * k + t;
                    st2(d1, (size_t)o, __ldg(sp), __ldg(sp + (size_t)C1 * k));
                }
            }
        } else {
            for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
                const int t = (int)(i % k);
                const int ci = (int)((i / k) % C1);
                const int co = (int)(i / ((long long)k * C1));
                const float v = j.src[i];
                if (d0) stf(d0, ((size_t)t * C0 + co) * C1 + ci, v);
                if (d1) stf(d1, ((size_t)t * C1 + ci) * C0 + co, v);
            }
        }
    }
}

// flat <-> scattered copies of many float tensors in one launch (gradient bucket of the data-parallel step)
__global__ void flat_copy_kernel(const pb_chunk_desc* __restrict__ chunks, float* __restrict__ flat, int to_flat, float scale) {
    const pb_chunk_desc c = chunks[blockIdx.y];
    float* t = c.ptr;
    float* f = flat + c.offset;
    const long long n4 = ((((uintptr_t)t | (uintptr_t)f) & 15) == 0) ? c.numel / 4 : 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (to_flat == 2) {                      // clear the range (ptr is not read)
        const long long z4 = (((uintptr_t)f & 15) == 0) ? c.numel / 4 : 0;
        for (long long i = tid; i < z4; i += stride) reinterpret_cast<float4*>(f)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (long long i = z4 * 4 + tid; i < c.numel; i += stride) f[i] = 0.f;
    } else if (to_flat) {
        for (long long i = tid; i < n4; i += stride) reinterpret_cast<float4*>(f)[i] = reinterpret_cast<const float4*>(t)[i];
        for (long long i = n4 * 4 + tid; i < c.numel; i += stride) f[i] = t[i];
    } else {
        for (long long i = tid; i < n4; i += stride) {
            float4 v = reinterpret_cast<const float4*>(f)[i];
            v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
            reinterpret_cast<float4*>(t)[i] = v;
        }
        for (long long i = n4 * 4 + tid; i < c.numel; i += stride) t[i] = f[i] * scale;
    }
}

extern "C" {

// Adam (torch.optim.Adam defaults of dummy_tests.py:127-130: no amsgrad, no weight decay unless asked) over many tensors in
// one launch: blockIdx.y walks a device table of <= 64k-element pieces.  Per element: 4 loads, 3 stores of 4 bytes.
__global__ void __launch_bounds__(256) adam_kernel(const pb_adam_desc* __restrict__ descs, float step_size, float beta1, float beta2,
                                                   float inv_sqrt_bc2, float eps, float weight_decay, float omb1, float omb2) {
    const pb_adam_desc d = descs[blockIdx.y];
    const bool vec = ((((uintptr_t)d.p | (uintptr_t)d.g | (uintptr_t)d.m | (uintptr_t)d.v) & 15) == 0);
    const long long n4 = vec ? d.numel / 4 : 0;
    const long long stride = (long long)gridDim.x * blockDim.x, tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    auto upd = [&](float& p, float g, float& m, float& v) {
        if (weight_decay != 0.f) g = fmaf(weight_decay, p, g);
        m = fmaf(beta1, m, omb1 * g);              // 1 - beta taken in double on the host, as torch does
        v = fmaf(beta2, v, omb2 * g * g);
        p -= step_size * m / (sqrtf(v) * inv_sqrt_bc2 + eps);
    };
    for (long long i = tid; i < n4; i += stride) {
        float4 p = reinterpret_cast<float4*>(d.p)[i], m = reinterpret_cast<float4*>(d.m)[i], v = reinterpret_cast<float4*>(d.v)[i];
        const float4 g = reinterpret_cast<const float4*>(d.g)[i];
        upd(p.x, g.x, m.x, v.x); upd(p.y, g.y, m.y, v.y); upd(p.z, g.z, m.z, v.z); upd(p.w, g.w, m.w, v.w);
        reinterpret_cast<float4*>(d.p)[i] = p; reinterpret_cast<float4*>(d.m)[i] = m; reinterpret_cast<float4*>(d.v)[i] = v;
    }
    for (long long i = n4 * 4 + tid; i < d.numel; i += stride) upd(d.p[i], d.g[i], d.m[i], d.v[i]);
}

int pb_adam_step(const pb_adam_desc* descs_dev, int n, float lr, double beta1, double beta2, float eps, float weight_decay, long long step,
                 void* stream) {
    if (n <= 0) return PB_OK;
    PB_REQUIRE(descs_dev != nullptr && step >= 1, PB_ERR_SHAPE, "adam_step: null table or step < 1");
    PB_REQUIRE(n <= 65535, PB_ERR_SHAPE, "adam_step: %d table rows exceed the 65535 one launch covers (gridDim.y); split the table", n);
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    adam_kernel<<<dim3(32, n), 256, 0, (cudaStream_t)stream>>>(descs_dev, (float)(lr / bc1), (float)beta1, (float)beta2, (float)(1.0 / sqrt(bc2)), eps,
                                                              weight_decay, (float)(1.0 - beta1), (float)(1.0 - beta2));
    pb_count_launch(1);
    return pb_check_launch("pb_adam_step");
}

int pb_flat_copy(const pb_chunk_desc* chunks_dev, int n, float* flat, int to_flat, float scale, void* stream) {
    if (n <= 0) return PB_OK;
    PB_REQUIRE(chunks_dev != nullptr && flat != nullptr, PB_ERR_SHAPE, "flat_copy: null argument");
    PB_REQUIRE(n <= 65535, PB_ERR_SHAPE, "flat_copy: %d table rows exceed the 65535 one launch covers (gridDim.y); split the table", n);
    flat_copy_kernel<<<dim3(8, n), 256, 0, (cudaStream_t)stream>>>(chunks_dev, flat, to_flat, scale);
    pb_count_launch(1);
    return pb_check_launch("pb_flat_copy");
}

int pb_stage_weights(const pb_stage_desc* jobs_dev, int n, int dt, void* stream) {
    if (n <= 0) return PB_OK;
    PB_REQUIRE(jobs_dev != nullptr, PB_ERR_SHAPE, "stage_weights: null job table");
    PB_REQUIRE(n <= 65535, PB_ERR_SHAPE, "stage_weights: %d jobs exceed the 65535 one launch covers (gridDim.y); split the table", n);
    dim3 grid(96, n);
    if (dt == PB_F32) stage_weights_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(jobs_dev);
    else if (dt == PB_BF16) stage_weights_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>(jobs_dev);
    else { pb_set_error("stage_weights: unknown dtype code %d", dt); return PB_ERR_UNSUPPORTED; }
    pb_count_launch(1);
    return pb_check_launch("pb_stage_weights");
}

int pb_version(void) { return 100; }
const char* pb_last_error(void) { return g_err; }
long long pb_launch_count(void) { return g_launches.load(); }

int pb_device_ok(int device) {
    cudaDeviceProp p;
    PB_CUDA(cudaGetDeviceProperties(&p, device));
    PB_REQUIRE(p.major == 10, PB_ERR_UNSUPPORTED,
               "device %d is sm_%d%d; libpb_b200 is built for sm_100a (B200) only and has no fallback", device,
               p.major, p.minor);
    return PB_OK;
}

int pb_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
    if (n <= 0) return PB_OK;
    PB_REQUIRE(((uintptr_t)src % 16 == 0) && ((uintptr_t)dst % 8 == 0), PB_ERR_SHAPE, "cast: misaligned pointers");
    const long long nthreads = (n + 3) / 4;
    cast_f32_bf16_kernel<<<ceil_div(nthreads, 256), 256, 0, (cudaStream_t)stream>>>(src, (bf16*)dst, n);
    pb_count_launch(1);
    return pb_check_launch("pb_cast_f32_to_bf16");
}

int pb_conv_weight_pack(const float* w, void* fwd_pack, void* bwd_pack, int C, int k, int dt, void* stream) {
    PB_REQUIRE(C > 0 && k > 0, PB_ERR_SHAPE, "conv_weight_pack: bad shape C=%d k=%d", C, k);
    const long long n = (long long)C * C * k;
    const int grid = min(ceil_div(n, 256), 1184);
    if (dt == PB_F32)
        conv_pack_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(w, (float*)fwd_pack, (float*)bwd_pack, C, k);
    else
        conv_pack_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>(w, (bf16*)fwd_pack, (bf16*)bwd_pack, C, k);
    pb_count_launch(1);
    return pb_check_launch("pb_conv_weight_pack");
}

size_t pb_workspace_bytes(const char* op, int M, int N, int K) {
    // split-K partials of the global-track GEMMs: at most 32 splits of an [M,N] (or [M,K]) float tile
    if (!strcmp(op, "linear_fwd")) return (size_t)32 * M * N * sizeof(float);
    if (!strcmp(op, "linear_bwd")) return (size_t)32 * M * K * sizeof(float);
    // split-K partial tiles of the tensor-core weight-gradient GEMMs: at most 148 tiles of 128 x 128 floats
    if (!strcmp(op, "wgrad")) return (size_t)148 * 3 * 128 * 128 * sizeof(float);     // three partial tiles per CTA (tc_wgrad_conv.cu)
    return 0;
}

}  // extern "C"

// ---- gradient exchange over NVLink / NVSwitch with the switch's in-fabric reduction (NVLS) ---------------------------------
// `mc` is the MULTICAST address of a symmetric buffer (every rank's copy of the flat gradient bucket mapped behind one
// address by torch.distributed._symmetric_memory).  Rank r owns elements [r*n/world, (r+1)*n/world): multimem.ld_reduce
// makes the switch fetch that range from every rank and add it (fp32), multimem.st writes the sum back into every rank's
// copy.  Two-shot all-reduce with each byte crossing each GPU's links once per direction; the callers bracket the launch
// with the symmetric-memory barrier (all ranks' gradients complete before, all ranges written after).
__device__ __forceinline__ float4 mm_ld_reduce(const float* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void mm_st(float* p, const float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// kInFlight reductions in flight per thread: the round trip through the switch is microseconds, so the loop is latency-bound.
// `ctas` encodes the launch shape: bits 0-15 = CTAs (0 -> 16), bits 16-27 = threads per CTA (0 -> 512), bits 28-30 = in-flight
// selector (0 -> 8, 1 -> 1, 2 -> 2, 3 -> 4).  The small shapes are for an exchange that runs BESIDE compute kernels: every
// outstanding reduction holds a load slot of its SM for the whole round trip, and CTAs of other kernels on that SM queue behind.
template <int kInFlight>
__global__ void __launch_bounds__(512) multimem_allreduce_kernel(float* mc, long long n4_begin, long long n4_end) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = n4_begin + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (kInFlight - 1) * stride < n4_end; i += kInFlight * stride) {
        float4 v[kInFlight];
#pragma unroll
        for (int u = 0; u < kInFlight; ++u) v[u] = mm_ld_reduce(mc + 4 * (i + u * stride));
#pragma unroll
        for (int u = 0; u < kInFlight; ++u) mm_st(mc + 4 * (i + u * stride), v[u]);
    }
    for (; i < n4_end; i += stride) mm_st(mc + 4 * i, mm_ld_reduce(mc + 4 * i));
}

extern "C" int pb_multimem_allreduce(float* mc_base, long long offset, long long numel, int rank, int world, int ctas, void* stream) {
    PB_REQUIRE(mc_base != nullptr && numel >= 0 && world >= 1 && rank >= 0 && rank < world, PB_ERR_SHAPE, "multimem_allreduce: bad arguments");
    PB_REQUIRE(offset % 4 == 0 && numel % 4 == 0 && ((uintptr_t)mc_base % 16) == 0, PB_ERR_SHAPE,
               "multimem_allreduce: offset and length must be multiples of 4 floats (16-byte vectors)");
    if (numel == 0) return PB_OK;
    const long long n4 = numel / 4, per = (n4 + world - 1) / world;
    const long long b = offset / 4 + (long long)rank * per, e = offset / 4 + ((long long)(rank + 1) * per < n4 ? (long long)(rank + 1) * per : n4);
    if (e <= b) return PB_OK;
    int threads = (ctas >> 16) & 0xfff;
    const int sel = (ctas >> 28) & 7;
    ctas &= 0xffff;
    if (ctas < 1) ctas = 16;
    if (threads < 32 || threads > 512) threads = 512;
    cudaStream_t st = (cudaStream_t)stream;
    switch (sel) {
        case 1: multimem_allreduce_kernel<1><<<ctas, threads, 0, st>>>(mc_base, b, e); break;
        case 2: multimem_allreduce_kernel<2><<<ctas, threads, 0, st>>>(mc_base, b, e); break;
        case 3: multimem_allreduce_kernel<4><<<ctas, threads, 0, st>>>(mc_base, b, e); break;
        default: multimem_allreduce_kernel<8><<<ctas, threads, 0, st>>>(mc_base, b, e); break;
    }
    pb_count_launch(1);
    return pb_check_launch("pb_multimem_allreduce");
}
