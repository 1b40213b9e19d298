/*
 * pb_b200.h -- C ABI of libpb_b200.so: ProteinBERT pretraining forward/backward on B200 (sm_100a).
 *
 * The reference (Aedelon/ProteinBERT-PyTorch-Replication) is pure Python over torch.nn and has no
 * native boundary of its own; every entry point below therefore cites the torch call sites of the
 * reference that it replaces (file:line relative to the reference root).  A binding needs nothing
 * but this header: plain device pointers, sizes, a dtype code and a CUDA stream.  No torch types.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless named host_*; nothing is allocated or freed here:
 *     outputs and workspaces are caller-owned (ask pb_workspace_bytes for the size).
 *   - layouts are row-major and channels-last: local track [B, L, C], global track [B, G].
 *   - dt selects the storage type of tensors marked `act` (activations) and `wt` (weights feeding
 *     the big local-track contractions): PB_F32 = float, fp32 SIMT arithmetic (parity mode);
 *     PB_BF16 = __nv_bfloat16 storage, tcgen05 bf16 tensor-core arithmetic with fp32 accumulation.
 *     Everything else (global track, LayerNorm affine, biases, gradients of parameters) is float;
 *     LayerNorm statistics and the loss are double.
 *   - stream is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises.
 *   - return value: 0 on success, PB_ERR_* otherwise, message via pb_last_error().  There is no
 *     CPU or library fallback: unsupported shapes are refused.
 */
#ifndef PB_B200_H
#define PB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB_OK 0
#define PB_ERR_SHAPE 1
#define PB_ERR_CUDA 2
#define PB_ERR_UNSUPPORTED 3
#define PB_ERR_WORKSPACE 4

#define PB_F32 0
#define PB_BF16 1

#define PB_ACT_NONE 0
#define PB_ACT_GELU 1
#define PB_ACT_SIGMOID 2

/* ---- library ------------------------------------------------------------------------------- */
int pb_version(void);
const char* pb_last_error(void);
/* 0 if `device` is a compute-capability-10.x GPU (the only target); PB_ERR_UNSUPPORTED otherwise. */
int pb_device_ok(int device);
/* number of kernel launches issued through this library since load (bench.py's gpu_launches). */
long long pb_launch_count(void);
/* 1 (default): PB_BF16 contractions run on the tcgen05 tensor-core engine where the tiling covers the shape
 * (C = 128 for the convolutions, C, G multiples of 128, L a multiple of 128); 0: force the fp32-accumulating
 * SIMT engine (same bf16 operands; used by the tests to cross-check the two engines); 4: tensor cores with the
 * first-generation (128-position, one TMA load per tap) conv kernel instead of the halo kernel.  Returns the old value. */
int pb_set_tensor_cores(int enabled);

/* ---- parameter staging ----------------------------------------------------------------------- */
/* float -> bf16 copy of a weight tensor (shadow weights of the bf16 mode). */
int pb_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream);
/* nn.Conv1d weight [Co, Ci, k] (modules.py:124-147) -> tap-major packs used by the conv kernels:
 *   fwd_pack [k][Co][Ci]  (forward / wgrad)  and  bwd_pack [k][Ci][Co] (data gradient), dtype dt. */
int pb_conv_weight_pack(const float* w, void* fwd_pack, void* bwd_pack, int C, int k, int dt, void* stream);

/* All per-step weight staging of the local track in ONE launch.  jobs_dev is a DEVICE array of n jobs:
 *   kind 0: src float [d0, d1] -> dst0[r*ld0 + c] (cast) and dst1[c*ld1 + r] (transposed cast); either may be NULL
 *   kind 1: src float [Co=d0, Ci=d1, k=d2] -> dst0 [k][Co][Ci] and dst1 [k][Ci][Co] (the two conv packs)
 * destinations have type dt.  Used for the conv packs, W2 / W2^T and Wv_all / Wv_all^T of every block. */
typedef struct pb_stage_desc {
    const float* src;
    void* dst0;
    void* dst1;
    int kind, d0, d1, d2, ld0, ld1, reserved;
} pb_stage_desc;
int pb_stage_weights(const pb_stage_desc* jobs_dev, int n, int dt, void* stream);

/* Gradient bucket of the data-parallel step: gather n float tensors into one flat
 * buffer (to_flat = 1), scatter the buffer back, multiplied by `scale` (to_flat = 0), or clear the listed ranges of the
 * buffer (to_flat = 2: the accumulated gradient slices, once per step; ptr is ignored), in one launch.  The
 * reference is single-process; this serves the one exchange step of the batch-sharded path (NCCL all-reduce). */
typedef struct pb_chunk_desc {
    float* ptr;
    long long offset; /* element offset inside the flat buffer */
    long long numel;
} pb_chunk_desc;
int pb_flat_copy(const pb_chunk_desc* chunks_dev, int n, float* flat, int to_flat, float scale, void* stream);

/* Gradient exchange with the NVSwitch's in-fabric reduction (no counterpart in the single-process reference).
 * mc_base: MULTICAST address of a symmetric buffer (torch.distributed._symmetric_memory: every rank's copy of the flat
 * bucket behind one address); the range [offset, offset + numel) floats is all-reduced (sum) in place on every rank: rank r
 * issues multimem.ld_reduce / multimem.st for its 1/world share.  Multiples of 4 floats.  The caller brackets the launch
 * with a cross-rank barrier on the same stream (before: every rank's gradients are complete; after: every share is
 * written).  ctas: launch shape -- bits 0-15 CTAs (0: 16), bits 16-27 threads per CTA (0: 512), bits 28-30 reductions in
 * flight per thread (0: 8, 1: 1, 2: 2, 3: 4); small shapes are for an exchange that runs beside compute kernels. */
int pb_multimem_allreduce(float* mc_base, long long offset, long long numel, int rank, int world, int ctas, void* stream);

/* ---- a20 / f-2: the optimizer step (utils.py:301 with Adam of dummy_tests.py:127-130), outside the metric's timed
 * region.  One launch updates every tensor of the table in place: m = b1 m + (1-b1) g, v = b2 v + (1-b2) g^2,
 * p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)  -- torch.optim.Adam's single-tensor formula (weight_decay adds
 * wd*p to g first; amsgrad/maximize not offered).  `step` is t >= 1; the betas are doubles because torch forms 1 - beta
 * in double before rounding to float.  Pieces of <= 64k elements fill the GPU.                                       */
typedef struct pb_adam_desc {
    float* p;
    const float* g;
    float* m;
    float* v;
    long long numel;
} pb_adam_desc;
int pb_adam_step(const pb_adam_desc* descs_dev, int n, float lr, double beta1, double beta2, float eps, float weight_decay,
                 long long step, void* stream);

/* ---- a2: token embedding (modules.py:249-253,300; backward = embedding_dense_backward) ------ */
int pb_embedding_fwd(const int64_t* tok, const float* table, void* out_act, long long n_tok, int C, int V,
                     int dt, void* stream);
int pb_embedding_bwd(const int64_t* tok, const void* dout_act, float* dtable, long long n_tok, int C, int V,
                     int dt, void* stream);

/* ---- a3/a7/a13/a14/a16: dense layers of the global track --------------------------------------
 * Y = act(X W^T + b), X [M,K], W [N,K] (nn.Linear layout), all float.  Replaces nn.Linear+GELU at
 * modules.py:255-262 (annotations -> global), :166-173 (global -> local), :175-182, :188-195 and
 * nn.Linear+Sigmoid at :286-293.  Z (pre-activation) is written when non-NULL (needed by the GELU
 * backward).  ws: split-K partials, pb_workspace_bytes("linear_fwd", M, N, K).                  */
int pb_linear_fwd(const float* X, const float* W, const float* bias, float* Z, float* Y, int M, int N, int K,
                  int act, void* ws, size_t ws_bytes, void* stream);
/* Same result as pb_linear_fwd, for callers that cannot provide (or create) a scratch buffer.  It has no
 * counterpart in the reference.  The split-K partial sums live in a 4 MiB device array that is part of the
 * library image (held by every GPU that loads the library, until exit), so the call allocates nothing.
 * Limit: M*N <= 32768 output elements per call (larger -> PB_ERR_UNSUPPORTED; use pb_linear_fwd).  The first
 * call on each device looks the array's address up (cudaGetSymbolAddress; under lazy module loading this may also load
 * the library's device code): make it before any stream capture on that device, with that device current -- `stream`
 * must belong to the current device.  All calls on a device share the array: calls that overlap on the GPU race --
 * serialise them.
 * Nothing in the array needs clearing between calls (every partial is rewritten before it is read).           */
int pb_linear_fwd_noalloc(const float* X, const float* W, const float* bias, float* Z, float* Y, int M, int N,
                          int K, int act, void* stream);
/* dZ = dY * act'(.) ; dX (+)= dZ W ; dW = dZ^T X ; db = sum_m dZ.  ZY is Z for GELU, Y for sigmoid.
 * dX may be NULL (annotations carry no gradient, modules.py:301); accumulate_dx adds into dX.
 * dZ_buf [M,N] is caller scratch.  ws as pb_workspace_bytes("linear_bwd", M, N, K).              */
int pb_linear_bwd(const float* dY, const float* ZY, const float* X, const float* W, float* dZ_buf, float* dX,
                  float* dW, float* db, int M, int N, int K, int act, int accumulate_dx, void* ws,
                  size_t ws_bytes, void* stream);

/* The same two operations on the tcgen05 tensor cores, for the layers whose weight matrix dwarfs the batch (the
 * annotation layers, modules.py:248-251 and 290-297: M = batch rows <= 128).  Same arguments and results; operands are
 * read as fp32, rounded to bf16 on the way into shared memory and accumulated in fp32 (the arithmetic of
 * precision = "bf16"; the fp32 entry points above keep full fp32).  ws is required: pb_linear_tc_workspace(M, N, K)
 * bytes cover both directions.  pb_linear_tc_ok says whether the shape is taken.  pb_linear_bwd_tc may be split in two calls
 * so that the weight gradient runs on another stream: first with dW = NULL (forms dZ_buf, db and dX), then with dY = NULL
 * and dX = NULL (dZ_buf is read as it stands; forms dW only).                                                             */
int pb_linear_tc_ok(int M, int N, int K);
size_t pb_linear_tc_workspace(int M, int N, int K);
int pb_linear_fwd_tc(const float* X, const float* W, const float* bias, float* Z, float* Y, int M, int N, int K,
                     int act, void* ws, size_t ws_bytes, void* stream);
int pb_linear_bwd_tc(const float* dY, const float* ZY, const float* X, const float* W, float* dZ_buf, float* dX,
                     float* dW, float* db, int M, int N, int K, int act, int accumulate_dx, void* ws,
                     size_t ws_bytes, void* stream);

/* y = LayerNorm_G(a + b + coef*c), coef = sum(w_param)/Kdim  (modules.py:183-186,196-199,223-229 and the
 * head mix of :87-92 under the reference's softmax axis).  b, c may be NULL.  Saves xhat and rstd. */
int pb_res_ln_rows_fwd(const float* a, const float* b, const float* c, const float* w_param, int Kdim,
                       const float* gamma, const float* beta, float* xhat, float* rstd, float* y, int B, int G,
                       void* stream);
/* dx = d(a) = d(b); d(c) = coef*dx (written to dc when non-NULL); dw_param[k] = (1/Kdim) sum dx*c. */
int pb_res_ln_rows_bwd(const float* dy, const float* gamma, const float* xhat, const float* rstd, const float* c,
                       const float* w_param, int Kdim, float* dx, float* dc, float* dgamma, float* dbeta,
                       float* dw_param, int B, int G, void* stream);

/* ---- a7/a13/a14 fused: the global side of one block in four launches (all float) ----------------------------
 * pre  : zg = g W1^T + b1, g2l = gelu(zg) (modules.py:208) and z3 = g W3^T + b3, gh = gelu(z3) (:221) in one GEMM pass.
 * post : g1 = LN(g + gh + coef*pool) (:92,:223-224), z4 = g1 W4^T + b4, g2 = LN(g1 + gelu(z4)) (:226-229).
 * The post kernels are cooperative launches (row phases and column phases separated by a grid barrier; at most
 * 148 CTAs).  pb_global_fused_ok says whether the shapes are covered (G, C multiples of 8, G <= 1024, G + C <=
 * 1184); otherwise compose pb_linear_* and pb_res_ln_rows_*. */
int pb_global_fused_ok(int B, int G, int C);
int pb_global_pre_fwd(const float* g, const float* W1, const float* b1, const float* W3, const float* b3, float* zg,
                      float* g2l, float* z3, float* gh, int B, int G, int C, void* stream);
/* dgh_is_dz3 = 1: the dgh argument already holds dgh * gelu'(z3) (the dz3 output of pb_global_post_bwd), which saves
 * every CTA the B x G GELU derivatives. */
int pb_global_pre_bwd(const float* dg2l, const float* dgh, const float* zg, const float* z3, const float* g,
                      const float* W1, const float* W3, float* dg, float* dW1, float* db1, float* dW3, float* db3,
                      int B, int G, int C, int dgh_is_dz3, float* dzg_out, void* stream);
/* Weight gradients off the critical path: pb_global_pre_bwd with dW1 = dW3 = NULL and pb_global_post_bwd with dW4 = NULL
 * skip their dW parts; the caller forms them as dW3 = dz3^T g, dW1 = dzg^T g (dzg_out [B,C] = dg2l * gelu'(zg), then
 * required) and dW4 = dz4^T g1 (dz4 = scratch3 + 2*B*G) with pb_matmul_tn, e.g. on a second stream.
 * out[n, k] = sum_m A[m, n] * B[m, k], all float, A [M,N], B [M,K], out [N,K]. */
int pb_matmul_tn(const float* A, const float* B, float* out, int M, int N, int K, void* stream);
int pb_global_post_fwd(const float* g, const float* gh, const float* pool, const float* Wp, int Kdim,
                       const float* gamma1, const float* beta1, const float* W4, const float* b4, const float* gamma2,
                       const float* beta2, float* g1, float* xhat1, float* rstd1, float* z4, float* xhat2, float* rstd2,
                       float* g2, int B, int G, void* stream);
/* scratch3 [3,B,G] is scratch; dx1 = d loss/d g (residual path) = d loss/d gh.  z3 / dz3 (both or neither): when given,
 * dz3 = dx1 * gelu'(z3) is written as well, for pb_global_pre_bwd(dgh = dz3, dgh_is_dz3 = 1). */
int pb_global_post_bwd(const float* dg2, const float* gamma2, const float* xhat2, const float* rstd2, const float* z4,
                       const float* g1, const float* W4, const float* gamma1, const float* xhat1, const float* rstd1,
                       const float* pool, const float* Wp, int Kdim, float* scratch3, float* dx1, float* dpool,
                       float* dWp, float* dgamma1, float* dbeta1, float* dW4, float* db4, float* dgamma2,
                       float* dbeta2, const float* z3, float* dz3, int B, int G, void* stream);

/* ---- a7/a13/a14 as one 16-CTA cluster launch per block boundary (csrc/gtrack.cu) --------------------------------
 * Forward unit = [post half of block i ; pre half of block i+1]:
 *   g1 = LN1(g + gh + (sum(Wp)/Kdim) pool), z4 = g1 W4^T + b4, g2 = LN2(g1 + gelu(z4))      (modules.py:92,223-229)
 *   zg = g2 W1^T + b1, g2l = gelu(zg), z3 = g2 W3^T + b3, gh_out = gelu(z3)                 (modules.py:208,221, next block)
 * W4 == NULL skips the post half (first block: the pre half then reads g_in); W1 == NULL skips the pre half (last
 * block).  All tensors float32; B <= 128, C <= G <= 512 (pb_gtrack_ok also asks the driver whether a 16-CTA cluster with
 * ~212 KB of shared memory per CTA can be scheduled); otherwise use pb_global_{pre,post}_*.
 * Backward unit = the mirror image [pre half of block i+1 ; post half of block i]:
 *   dzg = dg2l gelu'(zg); dg_total = dg_direct + dz3 W3 + dzg W1; db1, db3   (dz3_in given, or formed as dgh gelu'(z3n)
 *   into dz3_buf; dg_direct may be NULL)
 *   then with dg2 = dg_total (or dg_in when there is no pre half): LayerNorm-2 backward -> dv2, dz4 = dv2 gelu'(z4),
 *   dg1 = dv2 + dz4 W4, db4, dgamma2, dbeta2, LayerNorm-1 backward -> dv1 (gradient of g and of gh), dpool, dWp,
 *   dgamma1, dbeta1, and -- when z3 (this block's) is given -- dz3_out = dv1 gelu'(z3) for the next unit's dz3_in.
 * Weight gradients are left to the caller: dW4 = dz4^T g1, dW3 = dz3^T g, dW1 = dzg^T g (pb_matmul_tn).
 * dv2, dz4, dg1 are [B,G] scratch/outputs.  The batch is processed in groups of 16 rows, one cluster each; the groups
 * meet in the column sums, so db1, db3, db4, dgamma1, dbeta1, dgamma2, dbeta2 and dWp are ACCUMULATED: zero on entry. */
int pb_gtrack_ok(int B, int G, int C);
/* up to three weight-gradient products in one launch: out_j [N_j, K_j] = A_j^T X_j with A_j [M, N_j], X_j [M, K_j], M <= 128
 * (the batch); out_j == NULL skips product j.  Same results as three pb_matmul_tn calls. */
int pb_matmul_tn3(const float* A0, const float* X0, float* out0, int N0, int K0, const float* A1, const float* X1, float* out1,
                  int N1, int K1, const float* A2, const float* X2, float* out2, int N2, int K2, int M, void* stream);
int pb_gtrack_fwd(const float* g, const float* gh, const float* pool, const float* Wp, int Kdim, const float* gamma1,
                  const float* beta1, const float* W4, const float* b4, const float* gamma2, const float* beta2, float* g1,
                  float* xhat1, float* rstd1, float* z4, float* xhat2, float* rstd2, float* g2, const float* g_in,
                  const float* W1, const float* b1, const float* W3, const float* b3, float* zg, float* g2l, float* z3,
                  float* gh_out, int B, int G, int C, void* stream);
int pb_gtrack_bwd(const float* dg2l, const float* zg, const float* dz3_in, const float* dgh, const float* z3n,
                  const float* dg_direct, const float* W1, const float* W3, float* dz3_buf, float* dzg, float* db1, float* db3,
                  float* dg_total, const float* dg_in, const float* gamma2, const float* xhat2, const float* rstd2,
                  const float* z4, const float* W4, const float* gamma1, const float* xhat1, const float* rstd1,
                  const float* pool, const float* Wp, int Kdim, const float* z3, float* dv2, float* dz4, float* dg1, float* dv1,
                  float* dz3_out, float* dpool, float* dWp, float* dgamma1, float* dbeta1, float* db4, float* dgamma2,
                  float* dbeta2, int B, int G, int C, void* stream);

/* ---- a5/a6/a7: narrow + dilated Conv1d, GELU, 4-way add with the global broadcast ---------------
 * zn = conv(x, Wn, d=1)+bn, zw = conv(x, Ww, d=dil)+bw, u = x + gelu(zn) + gelu(zw) + g2l[b,:]
 * (modules.py:205-211), and stats[b] += (sum u, sum u^2) over (L,C) for the joint LayerNorm that
 * follows (:212).  stats must be zero on entry.  wn/ww are fwd packs from pb_conv_weight_pack. */
int pb_conv_pair_fwd(const void* x_act, const void* wn_pack, const void* ww_pack, const float* bn, const float* bw,
                     const float* g2l, void* zn_act, void* zw_act, void* u_act, double* stats, int B, int L, int C,
                     int ksize, int dil, int dt, void* stream);
/* backward of the above given du: dzn = du*gelu'(zn), dzw = du*gelu'(zw) (scratch, act type),
 * dx = du + dgrad_n(dzn) + dgrad_w(dzw) (+ dx_extra when non-NULL), dWn/dWw [Co,Ci,k] float, dbn/dbw,
 * dg2l[b,c] = sum_l du (dbn, dbw, dg2l are ACCUMULATED: zero on entry).  wn_bwd/ww_bwd are bwd packs.  ws: split-K partial tiles of the weight-gradient
 * GEMMs, pb_workspace_bytes("wgrad", 0, 0, 0) bytes (NULL or too small: atomics are used instead).   */
/* The three backward entry points of the local track (pb_conv_pair_bwd, pb_local_linear_bwd, pb_attn_pool_bwd) each have a
 * data-gradient half and a weight-gradient half that only the optimizer needs.  Passing NULL for the data-gradient
 * output (dx / dh / dh_out) runs the weight-gradient half alone, NULL for the weight gradient(s) (dWn+dWw / dW / dWvT) the
 * data-gradient half alone: a caller can put the weight gradients on a second stream, off the critical path of
 * backward.  In the weight-gradient-only call dzn/dzw, dz, dzv are INPUTS (written by the data-gradient call) and the
 * pre-activations z* must be NULL. */
int pb_conv_pair_bwd(const void* du_act, const void* zn_act, const void* zw_act, const void* x_act,
                     const void* wn_bwd, const void* ww_bwd, void* dzn_act, void* dzw_act, void* dx_act,
                     float* dWn, float* dWw, float* dbn, float* dbw, float* dg2l, int B, int L, int C, int ksize,
                     int dil, int dt, void* ws, size_t ws_bytes, void* stream);

/* ---- a8/a10: LayerNorm over (L,C) jointly, affine [L,C] (modules.py:148-151,161-164,212,217) --- */
int pb_ln_joint_stats(const void* u_act, double* stats, int B, int L, int C, int dt, void* stream);
int pb_ln_joint_apply(const void* u_act, const double* stats, const float* gamma, const float* beta, void* h_act,
                      int B, int L, int C, float eps, int dt, void* stream);
/* du, dgamma[L,C], dbeta[L,C] from dy; sums [B,2] double is caller scratch.  sums, dgamma and dbeta are
 * ACCUMULATED into: zero on entry. */
int pb_ln_joint_bwd(const void* dy_act, const void* u_act, const double* stats, const float* gamma, void* du_act,
                    float* dgamma, float* dbeta, double* sums, int B, int L, int C, float eps, int dt, void* stream);
/* The same LayerNorm backward fused with the GELU backward(s) that consume du in the block (modules.py:207-217 run in
 * reverse): additionally dz0 = du*gelu'(z0), db0[c] += sum dz0 and, when z1 is given, dz1/db1 likewise (z1, dz1, db1 all
 * NULL otherwise); dsample[b,c] += sum_l du when not NULL (the gradient of the broadcast global->local term).  db0, db1,
 * dsample, sums, dgamma, dbeta are ACCUMULATED into: zero on entry.  Needs (C/8) to divide 32 (pb_ln_joint_bwd_gelu_ok);
 * callers then pass z = NULL to pb_local_linear_bwd / zn = NULL to pb_conv_pair_bwd, which take dz as already computed. */
int pb_ln_joint_bwd_gelu_ok(int B, int L, int C);
int pb_ln_joint_bwd_gelu(const void* dy_act, const void* u_act, const double* stats, const float* gamma, const void* z0_act,
                         const void* z1_act, void* du_act, void* dz0_act, void* dz1_act, float* dgamma, float* dbeta,
                         float* dsample, float* db0, float* db1, double* sums, int B, int L, int C, float eps, int dt,
                         void* stream);

/* ---- a9: local Linear(C->C)+GELU + residual (modules.py:153-160,214-216) ------------------------
 * z = h W^T + b, u2 = h + gelu(z), stats2[b] += (sum, sumsq) of u2.  W is [C,C] of type wt.        */
int pb_local_linear_fwd(const void* h_act, const void* W_wt, const float* bias, void* z_act, void* u2_act,
                        double* stats2, int B, int L, int C, int dt, void* stream);
/* dz = du2*gelu'(z) (scratch), dh = du2 + dz W, dW = dz^T h, db = sum dz (db ACCUMULATED: zero on entry).  Wt_wt = W transposed ([C_in, C_out],
 * the K-major operand of the tensor-core data-gradient GEMM); may be NULL, which selects the SIMT engine.   */
int pb_local_linear_bwd(const void* du2_act, const void* z_act, const void* h_act, const void* W_wt, const void* Wt_wt,
                        void* dz_act,
                        void* dh_act, float* dW, float* db, int B, int L, int C, int dt, void* ws, size_t ws_bytes,
                        void* stream);

/* ---- a11/a12: global attention (modules.py:21-92) under the reference's softmax axis ------------
 * pool[b,g] = sum_l gelu(h[b,l,:] . Wv_all[:,g]); the layer output is (sum W_parameter / K) * pool
 * (Softmax(dim=1) over K identical query rows makes every weight exactly 1/K).
 * wvT is Wv_all transposed, [G, C], type wt.  pool is ACCUMULATED into: zero on entry.                                                     */
int pb_attn_pool_fwd(const void* h_act, const void* wvT_wt, float* pool, int B, int L, int C, int G, int dt,
                     void* stream);
/* dzv = dpool[b,g]*gelu'(h.Wv) (scratch [B*L,G] act), dh_out = dh_in + dzv Wv^T, dWvT [G,C] (NULL = skip).
 * wv_wt = Wv_all [C, G] (K-major operand of the tensor-core dh GEMM); NULL selects the SIMT engine.
 * head_dv > 0: dWvT receives one contiguous [C, head_dv] gradient per head (the layout of each head's
 * Wv_parameter, modules.py:36-39), stacked; head_dv = 0: the [G, C] matrix.                                  */
int pb_attn_pool_bwd(const void* h_act, const void* wvT_wt, const void* wv_wt, const float* dpool, const void* dh_in_act,
                     void* dzv_act, void* dh_out_act, float* dWvT, int head_dv, int B, int L, int C, int G, int dt, void* ws,
                     size_t ws_bytes, void* stream);

/* ---- a11/a12, op-faithful: GlobalAttention as the reference executes it, one fused launch per block (csrc/ops_attention.cu)
 * per head: q = tanh(g Wq) (the K query rows are identical, modules.py:51-53), kt = tanh(h Wk), s = q.kt^T / sqrt(K),
 * P = softmax(s) over `axis` (0: the repeated-query axis, the reference's nn.Softmax(dim=1), which makes every weight 1/K;
 * 1: the sequence positions -- the `softmax_axis="sequence"` mode), v = gelu(h Wv), out = P v;
 * att[b, head*Dv + d] = sum(W_parameter) * out[d] (:87-92).  Wq [H,G,K], Wk [H,C,K], Wv [H,C,Dv]: the heads' matrices stacked.
 * Scores, probabilities and values stay on chip; q_out [B,H,K], p_out [B,H,L], pool_out [B,G] are kept for backward.
 * Backward ACCUMULATES into dh [B,L,C] float, dg [B,G], dWq, dWk, dWv (zero on entry), writes dWp [K]; scratch = one float,
 * zero on entry.  fp32 SIMT arithmetic: this is the checker of the closed form (pb_attn_pool_*) and the only
 * implementation of axis 1. */
size_t pb_attention_smem(int L, int C, int K, int Dv, int backward);
int pb_attention_fwd(const void* h_act, const float* g, const float* Wq, const float* Wk, const float* Wv, const float* Wp, float* att,
                     float* q_out, float* p_out, float* pool_out, int B, int L, int C, int G, int K, int H, int axis, int dt,
                     void* stream);
int pb_attention_bwd(const void* h_act, const float* g, const float* Wq, const float* Wk, const float* Wv, const float* Wp,
                     const float* datt, const float* q, const float* p, const float* pool, float* dh, float* dg, float* dWq, float* dWk,
                     float* dWv, float* dWp, float* scratch, int B, int L, int C, int G, int K, int H, int axis, int dt, void* stream);

/* ---- a15: token head, softmax over the BATCH axis (modules.py:277-284; nn.Softmax() on 3-D) ----- */
int pb_token_head_fwd(const void* h_act, const float* Wo, const float* bo, float* p_local, int B, int L, int C,
                      int V, int dt, void* stream);
int pb_token_head_bwd(const float* dP, const float* p_local, const void* h_act, const float* Wo,
                      const double* grad_scale, void* dh_act, float* dWo, float* dbo, int B, int L, int C, int V,
                      int dt, void* stream);

/* Tensor-core variant of the token head (PB_BF16 activations; C, L multiples of 128; V <= 32).  wo_pad / woT_pad are
 * the output weight zero-padded to 128 vocabulary rows, [128, C] and its transpose [C, 128], type wt.  logits_ws
 * [B*L, 32] float and dlogit_ws [B*L, 128] wt are scratch; dWo_pad [128, C] float receives dWo in its first V rows;
 * dbo is ACCUMULATED (zero on entry); ws as pb_workspace_bytes("wgrad", 0, 0, 0).  pb_token_head_bwd_tc may be split in two
 * calls: first with dWo_pad = NULL (dlogit_ws, dbo, dh), then with dP = NULL and dh_act = NULL (reads dlogit_ws; dWo_pad). */
int pb_token_head_tc_ok(int B, int L, int C, int V);
int pb_token_head_fwd_tc(const void* h_act, const void* wo_pad, const float* bo, float* logits_ws, float* p_local, int B,
                         int L, int C, int V, void* stream);
int pb_token_head_bwd_tc(const float* dP, const float* p_local, const void* h_act, const void* woT_pad,
                         const double* grad_scale, void* dlogit_ws, void* dh_act, float* dWo_pad, float* dbo, int B, int L,
                         int C, int V, void* ws, size_t ws_bytes, void* stream);

/* ---- a17: the loss of utils.py:293-294 with CrossEntropyLoss/BCELoss(reduction='none') ----------
 * loss[0] = mean(CE(p_local over V, y_local) * w_local) + mean(BCE(p_global, y_global) * w_global)
 * (loss[1], loss[2] = the two terms), double.  When dP_local / dP_global are non-NULL they receive
 * d loss / d p for grad_output = 1.                                                                */
int pb_pretrain_loss(const float* p_local, const float* p_global, const int64_t* y_local, const float* y_global,
                     const double* w_local, const double* w_global, double* loss, float* dP_local,
                     float* dP_global, int B, int L, int V, int A, void* stream);
/* dZ = grad_scale * dP * p * (1 - p) for the sigmoid head (feeds pb_linear_bwd with PB_ACT_NONE). */
int pb_sigmoid_bwd(const float* dP, const float* p, const double* grad_scale, float* dZ, long long n, void* stream);

/* ---- a19: input corruption and padding (data_processing.py:30-183), bit-exact given the draws ----
 * seq_bytes: concatenated residue letters, seq_off [B+1]; draws in the reference's order:
 * crop_start [B] (ignored unless len+2 > L), tok_mask/tok_rand: concatenated per-token draws with
 * draw_off [B+1] (n_b = min(len_b+2, L) entries each), coin [B], keep/add [B,A] (0/1), ann [B,A].
 * Outputs: X_local,Y_local int64 [B,L]; X_global,Y_global float [B,A]; W_local double [B,L];
 * W_global double [B,A].                                                                           */
int pb_corrupt_pad(const uint8_t* seq_bytes, const int64_t* seq_off, const int64_t* crop_start,
                   const uint8_t* tok_mask, const uint8_t* tok_rand, const int64_t* draw_off, const float* coin,
                   const uint8_t* keep, const uint8_t* add, const uint8_t* ann, int64_t* x_local,
                   int64_t* y_local, double* w_local, float* x_global, float* y_global, double* w_global, int B,
                   int L, int A, void* stream);

/* ---- which engine a shape takes ------------------------------------------------------------------
 * op = "conv" (pb_conv_pair_*), "local_linear" (pb_local_linear_*), "attn_pool" (pb_attn_pool_*), all for PB_BF16
 * activations: 0 = the fp32-accumulating SIMT engine, 1 = the general tcgen05 kernels, 2 = the shape-specialised
 * tcgen05 kernels (halo-resident convolution / persistent pooling).  ksize/dil are read for "conv", G for "attn_pool".
 * Tests assert this before comparing engines, so a silent fall-back to SIMT cannot pass as a tensor-core result.
 * No counterpart in the reference (torch picks cuDNN / cuBLAS algorithms internally, modules.py:124-160). */
int pb_tc_path(const char* op, int B, int L, int C, int G, int ksize, int dil);

/* ---- f-3: the random draws of the data path made on the GPU (data_processing.py:78-83,100-105,125-142) ----------
 * Philox4x32-10 keyed by `seed`; `counter` distinguishes batches.  Fills the draw arguments of pb_corrupt_pad: crop_start
 * uniform in [0, len+2-L) where the sample is longer than L, tok_mask ~ Bernoulli(p_tok = 0.05), tok_rand uniform in
 * [3, 26), coin ~ U[0,1), keep ~ Bernoulli(p_keep = 0.75), add ~ Bernoulli(p_add = 1e-4).  Same distributions as the
 * reference's torch CPU draws, a different stream: statistically equivalent, not bit-identical. */
int pb_draw_batch(unsigned long long seed, unsigned long long counter, const int64_t* seq_off, const int64_t* draw_off,
                  int64_t* crop_start, uint8_t* tok_mask, uint8_t* tok_rand, float* coin, uint8_t* keep, uint8_t* add, int B,
                  int L, int A, float p_tok, float p_keep, float p_add, void* stream);

/* probe builds only (PB_NVCC_DEFINES=-DPB_GLOBAL_PROBE / -DPB_CONV2_PROBE): device buffer that receives phase
 * timestamps / cycle counters; NULL switches it off.  Process-global; the shipped library never writes to it. */
void pb_debug_buffer(void* device_buffer);

/* ---- sizing ------------------------------------------------------------------------------------ */
size_t pb_workspace_bytes(const char* op, int M, int N, int K);

#ifdef __cplusplus
}
#endif
#endif /* PB_B200_H */
