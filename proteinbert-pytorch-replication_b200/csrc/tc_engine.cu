// tcgen05 tensor-core engine -- placeholder until the kernels land: nothing qualifies, so the SIMT engine runs.
#include "tc_api.cuh"

bool pb_tc_conv_ok(int, int, int, int, int) { return false; }
bool pb_tc_linear_ok(int, int, int) { return false; }
bool pb_tc_attn_ok(int, int, int, int) { return false; }
bool pb_tc_ln_apply_ok(int, int) { return false; }

#define NOT_YET(name) pb_set_error(name ": tensor-core engine not available for this shape"); return PB_ERR_UNSUPPORTED
int pb_tc_conv_pair_fwd(const void*, const void*, const void*, const float*, const float*, const float*, void*, void*, void*,
                        double*, int, int, int, int, int, cudaStream_t) { NOT_YET("conv_pair_fwd"); }
int pb_tc_conv_pair_dgrad(const void*, const void*, const void*, const void*, const void*, void*, int, int, int, int, int,
                          cudaStream_t) { NOT_YET("conv_pair_dgrad"); }
int pb_tc_conv_wgrad(const void*, const void*, const void*, float*, float*, int, int, int, int, int, cudaStream_t) { NOT_YET("conv_wgrad"); }
int pb_tc_local_linear_fwd(const void*, const void*, const float*, void*, void*, double*, int, int, int, cudaStream_t) { NOT_YET("local_linear_fwd"); }
int pb_tc_local_linear_bwd(const void*, const void*, const void*, const void*, void*, float*, int, int, int, cudaStream_t) { NOT_YET("local_linear_bwd"); }
int pb_tc_attn_pool_fwd(const void*, const void*, float*, int, int, int, int, cudaStream_t) { NOT_YET("attn_pool_fwd"); }
int pb_tc_attn_pool_bwd(const void*, const void*, const float*, const void*, void*, void*, float*, int, int, int, int,
                        cudaStream_t) { NOT_YET("attn_pool_bwd"); }
int pb_tc_ln_apply(const void*, const double*, const float*, const float*, void*, int, int, int, float, cudaStream_t) { NOT_YET("ln_apply"); }
