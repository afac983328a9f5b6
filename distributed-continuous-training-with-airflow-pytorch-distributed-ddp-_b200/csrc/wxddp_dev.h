/* Developer aids of libwxddp.so.  NOT part of the drop-in boundary (include/wxddp.h): nothing in the product path
 * calls these; scripts/ and one GPU test of the tcgen05 primitives do.  The phase timers they feed exist only in builds
 * made with WX_NVCC_EXTRA=-DWX_TC_TIMERS=1. */
#pragma once
#include "wxddp.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Developer aid, not part of the reference-facing surface: when device_buf (int64 [grid][4]) is non-NULL the step
 * kernels record per CTA { smid, t_start, t_main_loop_end, t_exit } in globaltimer ns.  NULL switches it off. */
int wxdev_cta_times(wx_ctx* ctx, long long* device_buf);

/* Developer aid: runs the two tcgen05 (kind::tf32) GEMM shapes of the tensor-core step kernel on one 128-row tile
 * -- Z[128][64] = X[128][32] W[64][32]^T and G[64][32] = DH[128][64]^T X[128][32] -- with `terms` (1..3) terms of
 * the tf32 split, so descriptor encodings and TMEM layouts can be checked against a host matmul.  *status != 0 if
 * the MMA completion barrier timed out. */
int wxdev_tc_selftest(wx_ctx* ctx, const float* x, const float* w, const float* dh, int terms, float* z, float* g,
                   int* status, void* stream);

/* Developer aid: cycles for `reps` x 4 tcgen05.mma (kind::tf32) of one operand layout (variant 0..3, see
 * csrc/tc_selftest.cu); out = device int64[2] = { issue cycles, issue + completion cycles } of CTA 0. */
int wxdev_tc_mma_bench(wx_ctx* ctx, int variant, int n_dim, int reps, long long* out, void* stream);

#ifdef __cplusplus
}
#endif
