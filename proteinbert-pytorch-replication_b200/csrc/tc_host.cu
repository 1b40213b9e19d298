// Host side of the tensor-core engine: CUtensorMap construction through the driver entry point, the
// engine switch, and the vectorised bf16 LayerNorm-apply.
#include <atomic>
#include <stdlib.h>
#include <string.h>
#include <cudaTypedefs.h>
#include "tc_common.cuh"
#include "tc_api.cuh"
#include "../../include/pb_b200.h"

static std::atomic<int> g_tc_enabled{1};

static void* g_debug_ptr = nullptr;
void* pb_debug_ptr() { return g_debug_ptr; }
extern "C" void pb_debug_buffer(void* p) { g_debug_ptr = p; }

bool pb_tc_enabled() { return g_tc_enabled.load(std::memory_order_relaxed) != 0; }
int pb_tc_mode() { return g_tc_enabled.load(std::memory_order_relaxed); }

extern "C" int pb_set_tensor_cores(int enabled) { return g_tc_enabled.exchange(enabled < 0 ? 0 : enabled); }

// Encoded maps are kept in a small per-thread direct-mapped cache keyed by every argument of the encode call: an eager
// step re-creates the same ~200 maps per step (the conv kernels alone take nine each), and cuTensorMapEncodeTiled costs
// about a microsecond of host time per map.  Per thread because forward runs on the caller's thread and backward on
// autograd's worker: no lock on the launch path.  A map depends only on (address, shape, strides, box, swizzle), never on
// the tensor's contents, so a cached copy stays valid when the allocator hands the same address out again.
namespace {
struct TmapKey {
    const void* base; int rank, swizzle; uint64_t dims[3], strides[2]; uint32_t box[3];
    bool operator==(const TmapKey& o) const { return memcmp(this, &o, sizeof(TmapKey)) == 0; }
};
struct TmapSlot { TmapKey key; CUtensorMap map; bool valid; };
constexpr int kTmapSlots = 512;
}  // namespace

int pb_make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle_bytes) {
    static PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        PB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        PB_REQUIRE(fn != nullptr && qres == cudaDriverEntryPointSuccess, PB_ERR_CUDA, "cuTensorMapEncodeTiled not available");
        encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    }
    PB_REQUIRE(rank >= 2 && rank <= 3, PB_ERR_SHAPE, "tensor maps of rank 2 or 3 only");
    PB_REQUIRE(((uintptr_t)base % 16) == 0, PB_ERR_SHAPE, "tensor map base address must be 16-byte aligned");
    static thread_local TmapSlot* cache = nullptr;
    if (cache == nullptr) cache = (TmapSlot*)calloc(kTmapSlots, sizeof(TmapSlot));
    TmapKey key;
    memset(&key, 0, sizeof(key));
    key.base = base; key.rank = rank; key.swizzle = swizzle_bytes;
    for (int i = 0; i < rank; ++i) { key.dims[i] = dims[i]; key.box[i] = box[i]; }
    for (int i = 0; i + 1 < rank; ++i) key.strides[i] = strides_bytes[i];
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(key) / 8; ++i) { h ^= reinterpret_cast<const uint64_t*>(&key)[i]; h *= 1099511628211ull; }
    TmapSlot* slot = cache ? &cache[(h ^ (h >> 29)) % kTmapSlots] : nullptr;
    if (slot && slot->valid && slot->key == key) { *out = slot->map; return PB_OK; }
    cuuint64_t gdim[5];
    cuuint64_t gstr[4];
    cuuint32_t bdim[5], estr[5];
    for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bdim[i] = box[i]; estr[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) {
        PB_REQUIRE(strides_bytes[i] % 16 == 0, PB_ERR_SHAPE, "tensor map stride %d (%llu B) is not a multiple of 16", i,
                   (unsigned long long)strides_bytes[i]);
        gstr[i] = strides_bytes[i];
    }
    CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bdim, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE,
                        swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE),
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    PB_REQUIRE(r == CUDA_SUCCESS, PB_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    if (slot) { slot->key = key; slot->map = *out; slot->valid = true; }
    return PB_OK;
}

namespace {

// h = (u - mean) * rstd * gamma + beta, 8 bf16 per thread.  grid = (ceil(nv/256), sample groups): a thread owns one
// 8-wide (l,c) vector and walks its group's samples, so the [L,C] affine (8 bytes per element against 4 bytes of
// activation traffic) is read once per group instead of once per sample.
__global__ void ln_apply_bf16_kernel(const bf16* __restrict__ u, const double* __restrict__ stats, const float* __restrict__ gamma,
                                     const float* __restrict__ beta, bf16* __restrict__ h, long long n, float eps, int B,
                                     int per_group) {
    extern __shared__ float prm[];            // [per_group][2] mean, rstd
    pdl_wait();
    const int b0 = blockIdx.y * per_group;
    const int nb = min(per_group, B - b0);
    for (int k = threadIdx.x; k < nb; k += blockDim.x) {
        const LnStat st = ln_stat(stats, b0 + k, (double)n, eps);
        prm[2 * k] = st.mean; prm[2 * k + 1] = st.rstd;
    }
    __syncthreads();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n / 8) return;
    const float4 g0 = *reinterpret_cast<const float4*>(gamma + i * 8), g1 = *reinterpret_cast<const float4*>(gamma + i * 8 + 4);
    const float4 c0 = *reinterpret_cast<const float4*>(beta + i * 8), c1 = *reinterpret_cast<const float4*>(beta + i * 8 + 4);
    const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    const float bb[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll 4
    for (int k = 0; k < nb; ++k) {
        const size_t off = (size_t)(b0 + k) * n + i * 8;
        const uint4 uv = *reinterpret_cast<const uint4*>(u + off);
        const bf16* ub = reinterpret_cast<const bf16*>(&uv);
        const float mean = prm[2 * k], rstd = prm[2 * k + 1];
        __align__(16) bf16 o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = __float2bfloat16_rn((__bfloat162float(ub[j]) - mean) * rstd * gg[j] + bb[j]);
        *reinterpret_cast<uint4*>(h + off) = *reinterpret_cast<const uint4*>(o);
    }
}

}  // namespace

bool pb_tc_ln_apply_ok(int L, int C) { return ((long long)L * C) % 8 == 0; }

int pb_tc_ln_apply(const void* u, const double* stats, const float* gamma, const float* beta, void* h, int B, int L, int C,
                   float eps, cudaStream_t st) {
    const long long n = (long long)L * C, nv = n / 8;
    // enough sample groups for ~150k threads in flight, at most 16 samples per thread
    int groups = (int)((151552 + nv - 1) / nv);
    if (groups > B) groups = B;
    if (groups < (B + 15) / 16) groups = (B + 15) / 16;
    const int per_group = (B + groups - 1) / groups;
    groups = (B + per_group - 1) / per_group;
    dim3 grid((unsigned)ceil_div(nv, 256), groups);
    PB_CUDA(pb_launch_pdl(ln_apply_bf16_kernel, grid, dim3(256), 2 * per_group * sizeof(float), st, (const bf16*)u, stats, gamma, beta,
                          (bf16*)h, n, eps, B, per_group));
    pb_count_launch(1);
    return pb_check_launch("pb_ln_joint_apply(bf16)");
}
