// Host API of the tensor-core step kernel (csrc/tc_step_kernel.cuh): argument checks behind the C ABI.  The kernel is
// compiled three times (csrc/tc_step_inst.cu): for F = 20 and F = 5 with the feature count as a compile-time constant
// (BASELINE.json configs[1] and the reference's own table, preprocess.py:29), and once for any other F <= 31.
#include <string.h>

#include "mlp_kernels.cuh"

int wx_tc_launch_ft0(wx_ctx* ctx, wxk::GridArgs& a, cudaStream_t st, const char* who);
int wx_tc_launch_ft5(wx_ctx* ctx, wxk::GridArgs& a, cudaStream_t st, const char* who);
int wx_tc_launch_ft20(wx_ctx* ctx, wxk::GridArgs& a, cudaStream_t st, const char* who);

namespace {
using wxk::GridArgs;

int tc_launch(wx_ctx* ctx, GridArgs& a, cudaStream_t st, const char* who) {
    switch (ctx->desc.n_features) {
        case 5: return wx_tc_launch_ft5(ctx, a, st, who);
        case 20: return wx_tc_launch_ft20(ctx, a, st, who);
        default: return wx_tc_launch_ft0(ctx, a, st, who);
    }
}

int tc_check_common(wx_ctx* ctx, int& dropout_mode, const uint64_t* keep_mask, int64_t row0, int64_t batch, const char* who) {
    WX_CHECK_ARG(batch >= 1 && row0 >= 0, "%s: batch=%lld row0=%lld", who, (long long)batch, (long long)row0);
    WX_CHECK_ARG(dropout_mode == WX_DROPOUT_OFF || dropout_mode == WX_DROPOUT_MASK || dropout_mode == WX_DROPOUT_PHILOX,
                 "%s: dropout_mode=%d", who, dropout_mode);
    WX_CHECK_ARG(dropout_mode != WX_DROPOUT_MASK || keep_mask != nullptr, "%s: WX_DROPOUT_MASK needs keep_mask", who);
    if (ctx->desc.dropout_p == 0.0f) dropout_mode = WX_DROPOUT_OFF;
    return WX_OK;
}

}  // namespace

extern "C" {

int wx_tc_mlp_grad(wx_ctx* ctx, const float* params, const float* x, const int64_t* y, int64_t row0, int64_t batch,
                   int dropout_mode, const uint64_t* keep_mask, uint64_t seed, uint64_t step, float scale, float* out,
                   void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_tc_mlp_grad: ctx is NULL");
    WX_CHECK_ARG(params && x && y && out, "wx_tc_mlp_grad: NULL params/x/y/out");
    int rc = tc_check_common(ctx, dropout_mode, keep_mask, row0, batch, "wx_tc_mlp_grad");
    if (rc) return rc;
    GridArgs a;
    memset(&a, 0, sizeof(a));
    a.params = params;
    a.loader = wxk::LOAD_ROWMAJOR;
    a.x = x;
    a.y = y;
    a.row0 = row0;
    a.batch = batch;
    a.dropout_mode = dropout_mode;
    a.keep_mask = keep_mask;
    a.seed = seed;
    a.step = step;
    a.scale = scale;
    a.tail = wxk::TAIL_GRADS;
    a.out = out;
    return tc_launch(ctx, a, (cudaStream_t)stream, "wx_tc_mlp_grad");
}

int wx_tc_train_step(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state, const float* x, const int64_t* y,
                     int64_t row0, int64_t batch, int dropout_mode, const uint64_t* keep_mask, uint64_t seed,
                     float* loss_out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_tc_train_step: ctx is NULL");
    WX_CHECK_ARG(params && m && v && state && x && y && loss_out, "wx_tc_train_step: NULL pointer");
    int rc = tc_check_common(ctx, dropout_mode, keep_mask, row0, batch, "wx_tc_train_step");
    if (rc) return rc;
    GridArgs a;
    memset(&a, 0, sizeof(a));
    a.params = params;
    a.loader = wxk::LOAD_ROWMAJOR;
    a.x = x;
    a.y = y;
    a.row0 = row0;
    a.batch = batch;
    a.dropout_mode = dropout_mode;
    a.keep_mask = keep_mask;
    a.seed = seed;
    a.scale = 1.0f / float(batch);
    a.tail = wxk::TAIL_ADAM;
    a.p_rw = params;
    a.m = m;
    a.v = v;
    a.state = state;
    a.loss_out = loss_out;
    return tc_launch(ctx, a, (cudaStream_t)stream, "wx_tc_train_step");
}

int wx_tc_train_step_raw(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state, const void* const* cols,
                         const double* stats, int impute_mode, const int64_t* y, int64_t row0, int64_t batch,
                         int dropout_mode, const uint64_t* keep_mask, uint64_t seed, float* loss_out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_tc_train_step_raw: ctx is NULL");
    WX_CHECK_ARG(params && m && v && state && cols && stats && y && loss_out, "wx_tc_train_step_raw: NULL pointer");
    WX_CHECK_ARG(impute_mode == WX_IMPUTE_NONE || impute_mode == WX_IMPUTE_MEAN, "wx_tc_train_step_raw: impute_mode=%d",
                 impute_mode);
    int rc = tc_check_common(ctx, dropout_mode, keep_mask, row0, batch, "wx_tc_train_step_raw");
    if (rc) return rc;
    GridArgs a;
    memset(&a, 0, sizeof(a));
    const int f = ctx->desc.n_features;
    for (int j = 0; j < f; ++j) {
        WX_CHECK_ARG(cols[j] != nullptr, "wx_tc_train_step_raw: column %d is NULL", j);
        a.cols.col[j] = cols[j];
    }
    a.params = params;
    a.loader = wxk::LOAD_RAW_F32;
    a.stats = stats;
    a.impute = impute_mode;
    a.y = y;
    a.row0 = row0;
    a.batch = batch;
    a.dropout_mode = dropout_mode;
    a.keep_mask = keep_mask;
    a.seed = seed;
    a.scale = 1.0f / float(batch);
    a.tail = wxk::TAIL_ADAM;
    a.p_rw = params;
    a.m = m;
    a.v = v;
    a.state = state;
    a.loss_out = loss_out;
    return tc_launch(ctx, a, (cudaStream_t)stream, "wx_tc_train_step_raw");
}

/* Sticky protocol-error code of the tensor-core kernels (0 = none): a bounded mbarrier wait that expired. */
int wx_tc_last_status(wx_ctx* ctx, int* host_out) {
    WX_CHECK_ARG(ctx != nullptr && host_out != nullptr, "wx_tc_last_status: NULL argument");
    WX_CUDA(cudaSetDevice(ctx->device));
    WX_CUDA(cudaMemcpy(host_out, reinterpret_cast<int*>(ctx->scratch) + 8, sizeof(int), cudaMemcpyDeviceToHost));
    return WX_OK;
}

}  // extern "C"
