// Embedding, token head (batch-axis softmax), the pretraining loss and the sigmoid backward.
// Reference call sites: modules.py:249-253,277-293,300,304; utils.py:293-294; dummy_tests.py:132-133.
#include "common.cuh"
#include "../../include/pb_b200.h"
#include "tc_api.cuh"

namespace {

#define DISPATCH_DT(dt, ...)                                                   \
    do {                                                                       \
        if ((dt) == PB_F32) { using T = float; __VA_ARGS__; }                  \
        else if ((dt) == PB_BF16) { using T = bf16; __VA_ARGS__; }             \
        else { pb_set_error("unknown dtype code %d", (dt)); return PB_ERR_UNSUPPORTED; } \
    } while (0)

// ---- embedding --------------------------------------------------------------------------------
template <typename T>
__global__ void embedding_fwd_kernel(const int64_t* __restrict__ tok, const float* __restrict__ table, T* __restrict__ out,
                                     long long n_tok, int C, int V) {
    const long long total = n_tok * C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / C;
        const int c = (int)(i - r * C);
        const int64_t t = tok[r];
        stf(out, (size_t)i, (t >= 0 && t < V) ? table[(size_t)t * C + c] : 0.f);
    }
}
// C % 8 == 0: one thread per 8 channels of a row, 16-byte stores
__global__ void embedding_fwd_vec_kernel(const int64_t* __restrict__ tok, const float* __restrict__ table, bf16* __restrict__ out,
                                         long long n_tok, int C, int V) {
    const int vpr = C / 8;
    const long long total = n_tok * vpr;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / vpr;
        const int c = (int)(i - r * vpr) * 8;
        const int64_t t = tok[r];
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        if (t >= 0 && t < V) {
            a = __ldg(reinterpret_cast<const float4*>(table + (size_t)t * C + c));
            b = __ldg(reinterpret_cast<const float4*>(table + (size_t)t * C + c + 4));
        }
        uint4 o;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
        h[0] = __floats2bfloat162_rn(a.x, a.y); h[1] = __floats2bfloat162_rn(a.z, a.w);
        h[2] = __floats2bfloat162_rn(b.x, b.y); h[3] = __floats2bfloat162_rn(b.z, b.w);
        *reinterpret_cast<uint4*>(out + (size_t)r * C + c) = o;
    }
}

// shared-memory privatised scatter-add: each CTA owns a [V, C] float accumulator for its slice of rows
template <typename T>
__global__ void embedding_bwd_kernel(const int64_t* __restrict__ tok, const T* __restrict__ dout, float* __restrict__ dtable,
                                     long long n_tok, int C, int V) {
    extern __shared__ float acc[];           // [V*C]
    for (int i = threadIdx.x; i < V * C; i += blockDim.x) acc[i] = 0.f;
    __syncthreads();
    const long long per = (n_tok + gridDim.x - 1) / gridDim.x;
    const long long r0 = blockIdx.x * per, r1 = min(n_tok, r0 + per);
    const long long total = (r1 - r0) * C;
    for (long long i = threadIdx.x; i < total; i += blockDim.x) {
        const long long r = r0 + i / C;
        const int c = (int)(i % C);
        const int64_t t = tok[r];
        if (t >= 0 && t < V) atomicAdd(&acc[(int)t * C + c], ldf(dout, (size_t)r * C + c));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < V * C; i += blockDim.x) {
        const float v = acc[i];
        if (v != 0.f) atomicAdd(&dtable[i], v);
    }
}

// Warp-privatised variant (no atomics in the hot loop): every warp owns a full [V, C] accumulator in shared memory
// and walks its own rows; a row's token id is warp-uniform and the lanes split the channels, so each lane updates
// addresses nobody else touches.  The eight per-warp tables are summed once at the end.
template <typename T>
__global__ void embedding_bwd_warp_kernel(const int64_t* __restrict__ tok, const T* __restrict__ dout, float* __restrict__ dtable,
                                          long long n_tok, int C, int V) {
    extern __shared__ float acc[];           // [warps][V*C]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
    const int VC = V * C;
    for (int i = threadIdx.x; i < nwarp * VC; i += blockDim.x) acc[i] = 0.f;
    __syncthreads();
    float* mine = acc + (size_t)warp * VC;
    const long long gw = (long long)blockIdx.x * nwarp + warp, nw = (long long)gridDim.x * nwarp;
    // eight rows per trip: the token ids, then all their gradient values, are requested before anything is added, so a
    // warp's walk is a few round trips to memory instead of two per row
    constexpr int R = 8;
    for (long long r = gw; r < n_tok; r += R * nw) {
        int t[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const long long rr = r + j * nw;
            const int64_t tt = rr < n_tok ? tok[rr] : -1;
            t[j] = (tt >= 0 && tt < V) ? (int)tt : -1;
        }
        for (int c = lane; c < C; c += 32) {
            float v[R];
#pragma unroll
            for (int j = 0; j < R; ++j) v[j] = t[j] >= 0 ? ldf(dout, (size_t)(r + j * nw) * C + c) : 0.f;
#pragma unroll
            for (int j = 0; j < R; ++j)
                if (t[j] >= 0) mine[t[j] * C + c] += v[j];
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < VC; i += blockDim.x) {
        float v = 0.f;
        for (int w = 0; w < nwarp; ++w) v += acc[(size_t)w * VC + i];
        if (v != 0.f) atomicAdd(&dtable[i], v);
    }
}

// ---- token head ---------------------------------------------------------------------------------
// One CTA per sequence position l (grid-stride).  The softmax runs over the BATCH axis, so the CTA needs
// all B rows of its position: smem holds h[:, l, :] as float [B][C+1], Wo [V][C+1], logits/probs [B][V].
template <typename T>
__global__ void token_head_fwd_kernel(const T* __restrict__ h, const float* __restrict__ Wo, const float* __restrict__ bo,
                                      float* __restrict__ p_local, int B, int L, int C, int V) {
    extern __shared__ float sm[];
    const int ldc = C + 1;
    float* sh = sm;                          // [B][ldc]
    float* sw = sh + (size_t)B * ldc;        // [V][ldc]
    float* sp = sw + (size_t)V * ldc;        // [B][V]
    for (int i = threadIdx.x; i < V * C; i += blockDim.x) sw[(i / C) * ldc + (i % C)] = Wo[i];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    for (int l = blockIdx.x; l < L; l += gridDim.x) {
        __syncthreads();
        for (int i = threadIdx.x; i < B * C; i += blockDim.x) {
            const int b = i / C, c = i - b * C;
            sh[b * ldc + c] = ldf(h, ((size_t)b * L + l) * C + c);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < B * V; i += blockDim.x) {
            const int b = i / V, v = i - b * V;
            float a = bo[v];
            const float* hr = sh + b * ldc;
            const float* wr = sw + v * ldc;
            for (int c = 0; c < C; ++c) a = fmaf(hr[c], wr[c], a);
            sp[i] = a;
        }
        __syncthreads();
        // softmax over b for every vocabulary column v: one warp per column
        for (int v = warp; v < V; v += nwarp) {
            float mx = -INFINITY;
            for (int b = lane; b < B; b += 32) mx = fmaxf(mx, sp[b * V + v]);
            mx = warp_max(mx);
            float s = 0.f;
            for (int b = lane; b < B; b += 32) { const float e = expf(sp[b * V + v] - mx); sp[b * V + v] = e; s += e; }
            s = warp_sum(s);
            const float inv = 1.0f / s;
            for (int b = lane; b < B; b += 32) sp[b * V + v] *= inv;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < B * V; i += blockDim.x) {
            const int b = i / V, v = i - b * V;
            p_local[((size_t)b * L + l) * V + v] = sp[i];
        }
    }
}

// backward: dlogit[b,v] = s * p[b,v] * (dP[b,v] - sum_b' p[b',v] dP[b',v]);  dh = dlogit Wo;
// dWo[v,c] += sum_b dlogit[b,v] h[b,l,c];  dbo[v] += sum_b dlogit[b,v].
template <typename T>
__global__ void token_head_bwd_kernel(const float* __restrict__ dP, const float* __restrict__ p_local, const T* __restrict__ h,
                                      const float* __restrict__ Wo, const double* __restrict__ grad_scale, T* __restrict__ dh,
                                      float* __restrict__ dWo, float* __restrict__ dbo, int B, int L, int C, int V) {
    extern __shared__ float sm[];
    const int ldc = C + 1;
    float* sh = sm;                          // [B][ldc]
    float* sw = sh + (size_t)B * ldc;        // [V][ldc]
    float* sd = sw + (size_t)V * ldc;        // [B][V] dlogit
    float* sacc = sd + (size_t)B * V;        // [V][C] dWo partial
    float* sb = sacc + (size_t)V * C;        // [V] dbo partial
    const float gs = grad_scale ? (float)(*grad_scale) : 1.0f;
    for (int i = threadIdx.x; i < V * C; i += blockDim.x) { sw[(i / C) * ldc + (i % C)] = Wo[i]; sacc[i] = 0.f; }
    for (int i = threadIdx.x; i < V; i += blockDim.x) sb[i] = 0.f;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    for (int l = blockIdx.x; l < L; l += gridDim.x) {
        __syncthreads();
        for (int i = threadIdx.x; i < B * C; i += blockDim.x) {
            const int b = i / C, c = i - b * C;
            sh[b * ldc + c] = ldf(h, ((size_t)b * L + l) * C + c);
        }
        for (int v = warp; v < V; v += nwarp) {
            float dot = 0.f;
            for (int b = lane; b < B; b += 32) {
                const size_t gi = ((size_t)b * L + l) * V + v;
                dot += p_local[gi] * dP[gi];
            }
            dot = warp_sum(dot);
            float sbv = 0.f;
            for (int b = lane; b < B; b += 32) {
                const size_t gi = ((size_t)b * L + l) * V + v;
                const float d = gs * p_local[gi] * (dP[gi] - dot);
                sd[b * V + v] = d;
                sbv += d;
            }
            sbv = warp_sum(sbv);
            if (lane == 0) sb[v] += sbv;
        }
        __syncthreads();
        // dh[b,l,c] = sum_v dlogit[b,v] Wo[v,c]
        for (int i = threadIdx.x; i < B * C; i += blockDim.x) {
            const int b = i / C, c = i - b * C;
            float a = 0.f;
            for (int v = 0; v < V; ++v) a = fmaf(sd[b * V + v], sw[v * ldc + c], a);
            stf(dh, ((size_t)b * L + l) * C + c, a);
        }
        // dWo partial: thread owns (v,c) entries
        for (int i = threadIdx.x; i < V * C; i += blockDim.x) {
            const int v = i / C, c = i - v * C;
            float a = 0.f;
            for (int b = 0; b < B; ++b) a = fmaf(sd[b * V + v], sh[b * ldc + c], a);
            sacc[i] += a;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < V * C; i += blockDim.x) atomicAdd(&dWo[i], sacc[i]);
    for (int i = threadIdx.x; i < V; i += blockDim.x) atomicAdd(&dbo[i], sb[i]);
}

// ---- token head, tensor-core path: the GEMMs run in tc_gemm.cu; these are the batch-axis softmax around them ----
// logits32 [B*L, 32] fp32 (columns >= V unused) -> p_local [B, L, V].  One CTA per position l: 32 vocabulary lanes x 8
// batch lanes, each thread holds up to 8 samples of its (l, v) column in registers (all loads in flight at once); the
// batch-axis max / sum are combined through shared memory.  B <= 64 per register budget; larger batches loop.
constexpr int kTokBL = 8;            // batch lanes (warps) per CTA
constexpr int kTokPer = 8;           // samples per thread and trip
__global__ void __launch_bounds__(32 * kTokBL) softmax_batch_kernel(const float* __restrict__ logits32, const float* __restrict__ bo,
                                                                   float* __restrict__ p_local, int B, int L, int V) {
    __shared__ float sm_m[kTokBL][32], sm_s[kTokBL][32];
    const int l = blockIdx.x, v = threadIdx.x & 31, bl = threadIdx.x >> 5;
    const float bias = v < V ? bo[v] : 0.f;
    float mx = -INFINITY, s = 0.f;
    for (int b0 = bl; b0 < B; b0 += kTokBL * kTokPer) {          // online max / sum over this thread's samples
        float x[kTokPer];
#pragma unroll
        for (int i = 0; i < kTokPer; ++i) {
            const int b = b0 + i * kTokBL;
            x[i] = (b < B && v < V) ? logits32[((size_t)b * L + l) * 32 + v] + bias : -INFINITY;
        }
        float m2 = mx;
#pragma unroll
        for (int i = 0; i < kTokPer; ++i) m2 = fmaxf(m2, x[i]);
        if (m2 > -INFINITY) {
            s *= expf(mx - m2);
#pragma unroll
            for (int i = 0; i < kTokPer; ++i) s += expf(x[i] - m2);
            mx = m2;
        }
    }
    sm_m[bl][v] = mx; sm_s[bl][v] = s;
    __syncthreads();
    float M = -INFINITY;
#pragma unroll
    for (int i = 0; i < kTokBL; ++i) M = fmaxf(M, sm_m[i][v]);
    float S = 0.f;
#pragma unroll
    for (int i = 0; i < kTokBL; ++i) S += sm_m[i][v] > -INFINITY ? sm_s[i][v] * expf(sm_m[i][v] - M) : 0.f;
    if (v >= V) return;
    const float inv = 1.0f / S;
    for (int b = bl; b < B; b += kTokBL) {
        const float x = logits32[((size_t)b * L + l) * 32 + v] + bias;
        p_local[((size_t)b * L + l) * V + v] = expf(x - M) * inv;
    }
}

// dlogit[b,l,v] = s * p * (dP - sum_b' p dP) as bf16 into a [B*L, 128] buffer (columns >= V zero); dbo[v] += sum.
// Same layout: one CTA per position, 32 vocabulary lanes x 8 batch lanes.
__global__ void __launch_bounds__(32 * kTokBL) token_dlogit_kernel(const float* __restrict__ dP, const float* __restrict__ p_local,
                                                                  const double* __restrict__ gscale, bf16* __restrict__ dlogit_pad,
                                                                  float* __restrict__ dbo, int B, int L, int V) {
    __shared__ float sm_d[kTokBL][32];
    const int l = blockIdx.x, v = threadIdx.x & 31, bl = threadIdx.x >> 5;
    const float gs = gscale ? (float)(*gscale) : 1.0f;
    float dot = 0.f;
    if (v < V) {
#pragma unroll 8
        for (int b = bl; b < B; b += kTokBL) {
            const size_t gi = ((size_t)b * L + l) * V + v;
            dot += p_local[gi] * dP[gi];
        }
    }
    sm_d[bl][v] = dot;
    __syncthreads();
    dot = 0.f;
#pragma unroll
    for (int i = 0; i < kTokBL; ++i) dot += sm_d[i][v];
    __syncthreads();
    float sb = 0.f;
    const bf16 zero = __float2bfloat16_rn(0.f);
#pragma unroll 4
    for (int b = bl; b < B; b += kTokBL) {
        float d = 0.f;
        if (v < V) {
            const size_t gi = ((size_t)b * L + l) * V + v;
            d = gs * p_local[gi] * (dP[gi] - dot);
        }
        bf16* row = dlogit_pad + ((size_t)b * L + l) * 128;
        const bf16 db_ = __float2bfloat16_rn(d);
        row[v] = db_;
        row[v + 32] = zero; row[v + 64] = zero; row[v + 96] = zero;
        sb += __bfloat162float(db_);
    }
    sm_d[bl][v] = sb;
    __syncthreads();
    if (bl == 0 && v < V) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < kTokBL; ++i) t += sm_d[i][v];
        atomicAdd(&dbo[v], t);
    }
}

// ---- loss ---------------------------------------------------------------------------------------
// Token term: CE over the vocabulary axis of the PROBABILITIES (second softmax), times w_local, / (B*L).
// Annotation term: BCE with logs clamped at -100, times w_global, / (B*A).  fp64 accumulation.
__global__ void loss_kernel(const float* __restrict__ p_local, const float* __restrict__ p_global,
                            const int64_t* __restrict__ y_local, const float* __restrict__ y_global,
                            const double* __restrict__ w_local, const double* __restrict__ w_global, double* __restrict__ loss,
                            float* __restrict__ dP_local, float* __restrict__ dP_global, long long n_tok, int V, long long n_ann) {
    __shared__ double red[32];
    double acc_l = 0.0, acc_g = 0.0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const double inv_tok = 1.0 / (double)n_tok, inv_ann = 1.0 / (double)n_ann;
    if (V <= 32) {
        // one warp per token row, lane = vocabulary entry: coalesced row reads/writes, max and sum by shuffles, one expf
        const int lane = threadIdx.x & 31;
        const long long nwarp = stride >> 5;
        for (long long r = tid >> 5; r < n_tok; r += nwarp) {
            const float pv = lane < V ? p_local[(size_t)r * V + lane] : -INFINITY;
            const float mx = warp_max(pv);
            const float e = lane < V ? expf(pv - mx) : 0.f;
            const float ssum = warp_sum(e);
            const float lse = mx + logf(ssum);
            const int64_t y = y_local[r];
            const double w = w_local[r];
            const bool ok = (y >= 0 && y < V);
            if (ok && lane == (int)y) acc_l += (double)(lse - pv) * w;
            if (dP_local && lane < V) {
                const float wf = ok ? (float)(w * inv_tok) : 0.f;
                dP_local[(size_t)r * V + lane] = wf * (expf(pv - lse) - (lane == (int)y ? 1.f : 0.f));
            }
        }
    } else {
        for (long long r = tid; r < n_tok; r += stride) {
            const float* pr = p_local + (size_t)r * V;
            float mx = -INFINITY;
            for (int v = 0; v < V; ++v) mx = fmaxf(mx, pr[v]);
            float s = 0.f;
            for (int v = 0; v < V; ++v) s += expf(pr[v] - mx);
            const float lse = mx + logf(s);
            const int64_t y = y_local[r];
            const double w = w_local[r];
            const bool ok = (y >= 0 && y < V);
            if (ok) acc_l += (double)(lse - pr[y]) * w;
            if (dP_local) {
                const float wf = ok ? (float)(w * inv_tok) : 0.f;
                float* dr = dP_local + (size_t)r * V;
                for (int v = 0; v < V; ++v) dr[v] = wf * (expf(pr[v] - lse) - (v == (int)y ? 1.f : 0.f));
            }
        }
    }
    for (long long i = tid; i < n_ann; i += stride) {
        const float p = p_global[i], y = y_global[i];
        const double w = w_global[i];
        const float lp = fmaxf(logf(p), -100.f), l1p = fmaxf(logf(1.f - p), -100.f);   // (1 - input).log() as torch's BCELoss
        acc_g += (double)(-(y * lp + (1.f - y) * l1p)) * w;
        if (dP_global) {
            // torch binary_cross_entropy_backward: (p - y) / max((1 - p) * p, 1e-12)
            dP_global[i] = (float)(w * inv_ann) * (p - y) / fmaxf((1.f - p) * p, 1e-12f);
        }
    }
    acc_l = block_sum(acc_l, red) * inv_tok;
    acc_g = block_sum(acc_g, red) * inv_ann;
    if (threadIdx.x == 0) {
        atomicAdd(&loss[1], acc_l);
        atomicAdd(&loss[2], acc_g);
        atomicAdd(&loss[0], acc_l + acc_g);
    }
}

__global__ void sigmoid_bwd_kernel(const float* __restrict__ dP, const float* __restrict__ p, const double* __restrict__ gs,
                                   float* __restrict__ dZ, long long n) {
    const float s = gs ? (float)(*gs) : 1.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float pv = p[i];
        dZ[i] = s * dP[i] * pv * (1.f - pv);
    }
}

}  // namespace

extern "C" {

int pb_embedding_fwd(const int64_t* tok, const float* table, void* out, long long n_tok, int C, int V, int dt, void* stream) {
    PB_REQUIRE(n_tok > 0 && C > 0 && V > 0, PB_ERR_SHAPE, "embedding_fwd: bad shape");
    const int grid = (int)min((long long)ceil_div(n_tok * C, 256), 148LL * 16);
    if (dt == PB_BF16 && C % 8 == 0) {
        const int gv = (int)min((long long)ceil_div(n_tok * (C / 8), 256), 148LL * 16);
        embedding_fwd_vec_kernel<<<gv, 256, 0, (cudaStream_t)stream>>>(tok, table, (bf16*)out, n_tok, C, V);
    } else {
        DISPATCH_DT(dt, (embedding_fwd_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>(tok, table, (T*)out, n_tok, C, V)));
    }
    pb_count_launch(1);
    return pb_check_launch("pb_embedding_fwd");
}

int pb_embedding_bwd(const int64_t* tok, const void* dout, float* dtable, long long n_tok, int C, int V, int dt, void* stream) {
    PB_REQUIRE(n_tok > 0 && C > 0 && V > 0, PB_ERR_SHAPE, "embedding_bwd: bad shape");
    const size_t smem = (size_t)V * C * sizeof(float);
    PB_REQUIRE(smem <= 200 * 1024, PB_ERR_UNSUPPORTED, "embedding_bwd: table of %d x %d floats does not fit shared memory", V, C);
    cudaStream_t st = (cudaStream_t)stream;
    PB_CUDA(cudaMemsetAsync(dtable, 0, smem, st));
    const int grid = (int)min((long long)ceil_div(n_tok, 64), 148LL * 2);
    const size_t smem_w = smem * 8;          // one table per warp
    DISPATCH_DT(dt, {
        if (smem_w <= 200 * 1024) {
            if (smem_w > 48 * 1024) PB_CUDA(cudaFuncSetAttribute(embedding_bwd_warp_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_w));
            embedding_bwd_warp_kernel<T><<<min(grid, smem_w <= 110 * 1024 ? 296 : 148), 256, smem_w, st>>>(tok, (const T*)dout, dtable, n_tok, C, V);
        } else {
            if (smem > 48 * 1024) PB_CUDA(cudaFuncSetAttribute(embedding_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            embedding_bwd_kernel<T><<<grid, 256, smem, st>>>(tok, (const T*)dout, dtable, n_tok, C, V);
        }
    });
    pb_count_launch(1);
    return pb_check_launch("pb_embedding_bwd");
}

int pb_token_head_fwd(const void* h, const float* Wo, const float* bo, float* p_local, int B, int L, int C, int V, int dt,
                      void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0 && V > 0, PB_ERR_SHAPE, "token_head_fwd: bad shape");
    const size_t smem = ((size_t)B * (C + 1) + (size_t)V * (C + 1) + (size_t)B * V) * sizeof(float);
    PB_REQUIRE(smem <= 220 * 1024, PB_ERR_UNSUPPORTED,
               "token_head_fwd: the batch-axis softmax keeps all %d rows of a position on chip; B*C too large (%zu B)", B, smem);
    cudaStream_t st = (cudaStream_t)stream;
    DISPATCH_DT(dt, {
        if (smem > 48 * 1024) PB_CUDA(cudaFuncSetAttribute(token_head_fwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        token_head_fwd_kernel<T><<<min(L, 148 * 4), 256, smem, st>>>((const T*)h, Wo, bo, p_local, B, L, C, V);
    });
    pb_count_launch(1);
    return pb_check_launch("pb_token_head_fwd");
}

int pb_token_head_bwd(const float* dP, const float* p_local, const void* h, const float* Wo, const double* grad_scale, void* dh,
                      float* dWo, float* dbo, int B, int L, int C, int V, int dt, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && C > 0 && V > 0, PB_ERR_SHAPE, "token_head_bwd: bad shape");
    const size_t smem = ((size_t)B * (C + 1) + (size_t)V * (C + 1) + (size_t)B * V + (size_t)V * C + V) * sizeof(float);
    PB_REQUIRE(smem <= 220 * 1024, PB_ERR_UNSUPPORTED, "token_head_bwd: B*C too large for the on-chip batch softmax (%zu B)", smem);
    cudaStream_t st = (cudaStream_t)stream;
    PB_CUDA(cudaMemsetAsync(dWo, 0, (size_t)V * C * sizeof(float), st));
    PB_CUDA(cudaMemsetAsync(dbo, 0, (size_t)V * sizeof(float), st));
    DISPATCH_DT(dt, {
        if (smem > 48 * 1024) PB_CUDA(cudaFuncSetAttribute(token_head_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        token_head_bwd_kernel<T><<<min(L, 148 * 2), 256, smem, st>>>(dP, p_local, (const T*)h, Wo, grad_scale, (T*)dh, dWo, dbo, B, L, C, V);
    });
    pb_count_launch(1);
    return pb_check_launch("pb_token_head_bwd");
}

int pb_pretrain_loss(const float* p_local, const float* p_global, const int64_t* y_local, const float* y_global,
                     const double* w_local, const double* w_global, double* loss, float* dP_local, float* dP_global, int B,
                     int L, int V, int A, void* stream) {
    PB_REQUIRE(B > 0 && L > 0 && V > 0 && A > 0, PB_ERR_SHAPE, "pretrain_loss: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    PB_CUDA(cudaMemsetAsync(loss, 0, 3 * sizeof(double), st));
    const long long n_tok = (long long)B * L, n_ann = (long long)B * A;
    // one resident wave of fat CTAs: every CTA ends with three same-address double atomics, so there are few of them
    const int grid = (int)min((long long)ceil_div(n_ann > n_tok ? n_ann : n_tok, 512), 148LL * 2);
    loss_kernel<<<grid, 512, 0, st>>>(p_local, p_global, y_local, y_global, w_local, w_global, loss, dP_local, dP_global, n_tok,
                                      V, n_ann);
    pb_count_launch(1);
    return pb_check_launch("pb_pretrain_loss");
}

int pb_sigmoid_bwd(const float* dP, const float* p, const double* grad_scale, float* dZ, long long n, void* stream) {
    if (n <= 0) return PB_OK;
    sigmoid_bwd_kernel<<<(int)min((long long)ceil_div(n, 256), 148LL * 8), 256, 0, (cudaStream_t)stream>>>(dP, p, grad_scale, dZ, n);
    pb_count_launch(1);
    return pb_check_launch("pb_sigmoid_bwd");
}

}  // extern "C"

extern "C" {

int pb_token_head_tc_ok(int B, int L, int C, int V) { return pb_tc_token_ok(B, L, C, V) ? 1 : 0; }

int pb_token_head_fwd_tc(const void* h, const void* wo_pad, const float* bo, float* logits_ws, float* p_local, int B, int L, int C,
                         int V, void* stream) {
    PB_REQUIRE(pb_tc_token_ok(B, L, C, V), PB_ERR_UNSUPPORTED, "token_head_fwd_tc: shape not covered (C %% 128, L %% 128, V <= 32)");
    cudaStream_t st = (cudaStream_t)stream;
    int rc = pb_tc_token_logits(h, wo_pad, logits_ws, B, L, C, st);
    if (rc) return rc;
    softmax_batch_kernel<<<L, 32 * kTokBL, 0, st>>>(logits_ws, bo, p_local, B, L, V);
    pb_count_launch(1);
    return pb_check_launch("pb_token_head_fwd_tc");
}

int pb_token_head_bwd_tc(const float* dP, const float* p_local, const void* h, const void* woT_pad, const double* grad_scale,
                         void* dlogit_ws, void* dh, float* dWo_pad, float* dbo, int B, int L, int C, int V, void* ws,
                         size_t ws_bytes, void* stream) {
    PB_REQUIRE(pb_tc_token_ok(B, L, C, V), PB_ERR_UNSUPPORTED, "token_head_bwd_tc: shape not covered");
    cudaStream_t st = (cudaStream_t)stream;
    if (dP != nullptr) {       // dP == NULL: dlogit_ws already holds the logit gradients of an earlier call (and dbo is done)
        token_dlogit_kernel<<<L, 32 * kTokBL, 0, st>>>(dP, p_local, grad_scale, (bf16*)dlogit_ws, dbo, B, L, V);
        pb_count_launch(1);
        int rc = pb_check_launch("pb_token_head_bwd_tc(dlogit)");
        if (rc) return rc;
    }
    return pb_tc_token_bwd(dlogit_ws, woT_pad, h, dh, dWo_pad, B, L, C, ws, ws_bytes, st);
}

}  // extern "C"
