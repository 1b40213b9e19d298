// Internal interface between the op entry points and the tcgen05 tensor-core engine (tc_*.cu).
// *_ok() says whether the tensor-core tiling covers a shape; if not the fp32 SIMT engine runs instead
// (same results up to summation order: both use bf16 operands with fp32 accumulation in PB_BF16 mode).
#pragma once
#include "common.cuh"

bool pb_tc_enabled();
int pb_tc_mode();      // value given to pb_set_tensor_cores: 0 off, 1 default, 4 = force the first-generation conv kernel
bool pb_tc_conv_ok(int B, int L, int C, int ksize, int dil);
bool pb_tc_linear_ok(int B, int L, int C);
bool pb_tc_attn_ok(int B, int L, int C, int G);
bool pb_tc_ln_apply_ok(int L, int C);

int pb_tc_conv_pair_fwd(const void* x, const void* wn, const void* ww, const float* bn, const float* bw, const float* g2l,
                        void* zn, void* zw, void* u, double* stats, int B, int L, int C, int ksize, int dil, cudaStream_t st);
int pb_tc_conv_pair_dgrad(const void* du, const void* dzn, const void* dzw, const void* wn_bwd, const void* ww_bwd, void* dx,
                          int B, int L, int C, int ksize, int dil, cudaStream_t st);
int pb_tc_conv_wgrad(const void* dzn, const void* dzw, const void* x, float* dWn, float* dWw, int B, int L, int C, int ksize,
                     int dil, void* ws, size_t ws_bytes, cudaStream_t st);
bool pb_tc_conv_wgrad3_ok(int B, int L, int C, int ksize, int dil, size_t ws_bytes);
int pb_tc_conv_wgrad3(const void* dzn, const void* dzw, const void* x, float* dWn, float* dWw, int B, int L, int ksize, int dil,
                      void* ws, cudaStream_t st);
int pb_tc_local_linear_fwd(const void* h, const void* W, const float* bias, void* z, void* u2, double* stats2, int B, int L,
                           int C, cudaStream_t st);
int pb_tc_local_linear_bwd(const void* du2, const void* dz, const void* h, const void* Wt, void* dh, float* dW, int B, int L,
                           int C, void* ws, size_t ws_bytes, cudaStream_t st);
int pb_tc_attn_pool_fwd(const void* h, const void* wvT, float* pool, int B, int L, int C, int G, cudaStream_t st);
int pb_tc_attn_pool_bwd(const void* h, const void* wvT, const void* wv, const float* dpool, const void* dh_in, void* dzv, void* dh_out,
                        float* dWvT, int head_dv, int B, int L, int C, int G, void* ws, size_t ws_bytes, cudaStream_t st);
int pb_tc_ln_apply(const void* u, const double* stats, const float* gamma, const float* beta, void* h, int B, int L, int C,
                   float eps, cudaStream_t st);

// persistent pooling GEMMs (tc_pool.cu): C = 128, G <= 512
bool pb_tc_pool_ok(int B, int L, int C, int G);
int pb_tc_pool_fwd(const void* h, const void* wvT, float* pool, int B, int L, int G, cudaStream_t st);
int pb_tc_pool_dz(const void* h, const void* wvT, const float* dpool, void* dzv, int B, int L, int G, cudaStream_t st);

bool pb_tc_conv2_ok(int B, int L, int C, int ksize, int dil);
int pb_tc_conv2_fwd(const void* x, const void* wn, const void* ww, const float* bn, const float* bw, const float* g2l, void* zn,
                    void* zw, void* u, double* stats, int B, int L, int C, int ksize, int dil, int bo_mode, cudaStream_t st);
int pb_tc_conv2_dgrad(const void* du, const void* dzn, const void* dzw, const void* wn_bwd, const void* ww_bwd, void* dx, int B,
                      int L, int C, int ksize, int dil, int bo_mode, cudaStream_t st);
void* pb_debug_ptr();
bool pb_tc_token_ok(int B, int L, int C, int V);
int pb_tc_token_logits(const void* h, const void* wo_pad, float* logits32, int B, int L, int C, cudaStream_t st);
int pb_tc_token_bwd(const void* dlogit_pad, const void* woT_pad, const void* h, void* dh, float* dWo_pad, int B, int L, int C,
                    void* ws, size_t ws_bytes, cudaStream_t st);

