// Host side of K2/K3/K5: argument checks, grid sizing and launches behind the C ABI (include/wxddp.h).
#include <mutex>

#include "mlp_kernels.cuh"

namespace {
using namespace wxk;

grid_kernel_t pick_grid_kernel(int f, int mode, int loader, int* smem) {
    switch ((f - 1) / 8) {
        case 0: return pick_grid_kernel_g0(f, mode, loader, smem);
        case 1: return pick_grid_kernel_g1(f, mode, loader, smem);
        case 2: return pick_grid_kernel_g2(f, mode, loader, smem);
        case 3: return pick_grid_kernel_g3(f, mode, loader, smem);
        default: return nullptr;
    }
}
small_kernel_t pick_small_kernel(int f, int* smem) {
    switch ((f - 1) / 8) {
        case 0: return pick_small_kernel_g0(f, smem);
        case 1: return pick_small_kernel_g1(f, smem);
        case 2: return pick_small_kernel_g2(f, smem);
        case 3: return pick_small_kernel_g3(f, smem);
        default: return nullptr;
    }
}

small_kernel_t pick_resident_kernel(int f, int* smem) {
    switch ((f - 1) / 8) {
        case 0: return pick_resident_kernel_g0(f, smem);
        case 1: return pick_resident_kernel_g1(f, smem);
        case 2: return pick_resident_kernel_g2(f, smem);
        case 3: return pick_resident_kernel_g3(f, smem);
        default: return nullptr;
    }
}

int launch_grid(wx_ctx* ctx, GridArgs& a, int mode, cudaStream_t st, const char* who) {
    const int f = ctx->desc.n_features;
    int smem = 0;
    grid_kernel_t k = pick_grid_kernel(f, mode, a.loader, &smem);
    WX_CHECK_ARG(k != nullptr, "%s: no kernel for F=%d", who, f);
    WX_CUDA(cudaSetDevice(ctx->device));
    WX_CUDA(cudaFuncSetAttribute(reinterpret_cast<const void*>(k), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int occ = 0;
    WX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, reinterpret_cast<const void*>(k), THREADS, smem));
    if (occ < 1) occ = 1;
    const int64_t nchunks = (a.batch + R - 1) / R;
    int64_t grid = (nchunks + WARPS - 1) / WARPS;
    const int64_t cap = int64_t(ctx->num_sms) * occ;
    if (grid > cap) grid = cap;
    if (grid > WX_MAX_GRID_PARTS) grid = WX_MAX_GRID_PARTS;
    if (grid < 1) grid = 1;
    const size_t slab = (size_t(64) * f + 64 + 128 + 2 + 1 + 3) & ~size_t(3);
    // one partial slab per CTA: never launch more CTAs than the scratch area holds slabs for (the kernels walk their
    // chunks grid-stride, so a smaller grid computes the same sums) -- this is what bounds the grid of the fixed
    // 1.5 MiB array behind wx_mlp_grad_static
    if (ctx->scratch_bytes > WX_SCRATCH_PART_OFF) {
        const int64_t fit = int64_t((ctx->scratch_bytes - WX_SCRATCH_PART_OFF) / (mode == MODE_TRAIN ? slab * 4 : 16));
        if (fit >= 1 && grid > fit) grid = fit;
    }
    const size_t need = mode == MODE_TRAIN ? size_t(grid) * slab * 4 : size_t(grid) * 16;
    if (WX_SCRATCH_PART_OFF + need > ctx->scratch_bytes) {
        wx_set_error("%s: scratch too small (%zu needed)", who, need);
        return WX_ENOMEM;
    }
    a.part = reinterpret_cast<float*>(ctx->scratch + WX_SCRATCH_PART_OFF);
    a.ticket = reinterpret_cast<unsigned int*>(ctx->scratch) + (mode == MODE_TRAIN ? 1 : 2);
    a.desc = ctx->desc;
    a.dbg = ctx->debug_times;
    void* args[] = {&a};
    WX_CUDA(cudaLaunchKernel(reinterpret_cast<const void*>(k), dim3((unsigned)grid), dim3(THREADS), args, smem, st));
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

int check_dropout(const wx_ctx* ctx, int& mode, const uint64_t* keep_mask, const char* who) {
    if (ctx->desc.dropout_p == 0.0f) mode = WX_DROPOUT_OFF;   // a model built without dropout never drops
    WX_CHECK_ARG(mode == WX_DROPOUT_OFF || mode == WX_DROPOUT_MASK || mode == WX_DROPOUT_PHILOX, "%s: dropout_mode=%d",
                 who, mode);
    WX_CHECK_ARG(mode != WX_DROPOUT_MASK || keep_mask != nullptr, "%s: WX_DROPOUT_MASK needs keep_mask", who);
    return WX_OK;
}

void zero_args(GridArgs& a) { memset(&a, 0, sizeof(a)); }

}  // namespace

extern "C" {

int wx_mlp_grad(wx_ctx* ctx, const float* params, const float* x, const int64_t* y, const int64_t* row_idx,
                int64_t row0, int64_t batch, int dropout_mode, const uint64_t* keep_mask, uint64_t seed, uint64_t step,
                float scale, float* out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_mlp_grad: ctx is NULL");
    WX_CHECK_ARG(params && x && y && out, "wx_mlp_grad: NULL params/x/y/out");
    WX_CHECK_ARG(batch >= 1, "wx_mlp_grad: batch=%lld must be >= 1", (long long)batch);
    WX_CHECK_ARG(row_idx != nullptr || row0 >= 0, "wx_mlp_grad: row0 < 0");
    int rc = check_dropout(ctx, dropout_mode, keep_mask, "wx_mlp_grad");
    if (rc) return rc;
    GridArgs a;
    zero_args(a);
    a.params = params;
    a.loader = LOAD_ROWMAJOR;
    a.x = x;
    a.y = y;
    a.row_idx = row_idx;
    a.row0 = row0;
    a.batch = batch;
    a.dropout_mode = dropout_mode;
    a.keep_mask = keep_mask;
    a.seed = seed;
    a.step = step;
    a.scale = scale;
    a.tail = TAIL_GRADS;
    a.out = out;
    return launch_grid(ctx, a, MODE_TRAIN, (cudaStream_t)stream, "wx_mlp_grad");
}

int wx_adam_step(wx_ctx* ctx, float* params, const float* grads, float* m, float* v, int64_t n, int64_t step,
                 void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_adam_step: ctx is NULL");
    WX_CHECK_ARG(params && grads && m && v, "wx_adam_step: NULL pointer");
    WX_CHECK_ARG(n >= 0 && step >= 1, "wx_adam_step: n=%lld step=%lld", (long long)n, (long long)step);
    if (n == 0) return WX_OK;
    WX_CUDA(cudaSetDevice(ctx->device));
    const wx_model_desc& d = ctx->desc;
    AdamScalars a;
    a.sgd = d.optimizer == WX_OPT_SGD;
    const double bc1 = a.sgd ? 1.0 : 1.0 - pow(d.beta1, double(step));
    const double bc2 = 1.0 - pow(d.beta2, double(step));
    a.w_lerp = float(1.0 - d.beta1);
    a.beta2 = float(d.beta2);
    a.one_m_beta2 = float(1.0 - d.beta2);
    a.neg_step = float(-(d.lr / bc1));
    a.bc2_sqrt = float(sqrt(bc2));
    a.eps = float(d.eps);
    const uintptr_t al = reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) |
                         reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v);
    const int vec_ok = (al & 15) == 0;
    const int64_t work = vec_ok ? ((n + 3) >> 2) : n;
    int64_t blocks = (work + 255) / 256;
    const int64_t cap = int64_t(ctx->num_sms) * 8;
    if (blocks > cap) blocks = cap;
    adam_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(params, grads, m, v, n, a, vec_ok);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

int wx_train_step(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state, const float* x, const int64_t* y,
                  const int64_t* row_idx, int64_t row0, int64_t batch, int dropout_mode, const uint64_t* keep_mask,
                  uint64_t seed, float* loss_out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_train_step: ctx is NULL");
    WX_CHECK_ARG(params && m && v && state && x && y && loss_out, "wx_train_step: NULL pointer");
    WX_CHECK_ARG(batch >= 1, "wx_train_step: batch=%lld must be >= 1", (long long)batch);
    WX_CHECK_ARG(row_idx != nullptr || row0 >= 0, "wx_train_step: row0 < 0");
    int rc = check_dropout(ctx, dropout_mode, keep_mask, "wx_train_step");
    if (rc) return rc;
    GridArgs a;
    zero_args(a);
    a.params = params;
    a.loader = LOAD_ROWMAJOR;
    a.x = x;
    a.y = y;
    a.row_idx = row_idx;
    a.row0 = row0;
    a.batch = batch;
    a.dropout_mode = dropout_mode;
    a.keep_mask = keep_mask;
    a.seed = seed;
    a.scale = 1.0f / float(batch);
    a.tail = TAIL_ADAM;
    a.p_rw = params;
    a.m = m;
    a.v = v;
    a.state = state;
    a.loss_out = loss_out;
    return launch_grid(ctx, a, MODE_TRAIN, (cudaStream_t)stream, "wx_train_step");
}

int wx_train_step_raw(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state, const void* const* cols,
                      int dtype, const double* stats, int impute_mode, const int64_t* y, int64_t row0, int64_t batch,
                      int dropout_mode, const uint64_t* keep_mask, uint64_t seed, float* loss_out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_train_step_raw: ctx is NULL");
    WX_CHECK_ARG(params && m && v && state && cols && stats && y && loss_out, "wx_train_step_raw: NULL pointer");
    WX_CHECK_ARG(batch >= 1 && row0 >= 0, "wx_train_step_raw: batch=%lld row0=%lld", (long long)batch, (long long)row0);
    WX_CHECK_ARG(dtype == WX_F32 || dtype == WX_F64, "wx_train_step_raw: dtype=%d", dtype);
    WX_CHECK_ARG(impute_mode == WX_IMPUTE_NONE || impute_mode == WX_IMPUTE_MEAN, "wx_train_step_raw: impute_mode=%d",
                 impute_mode);
    int rc = check_dropout(ctx, dropout_mode, keep_mask, "wx_train_step_raw");
    if (rc) return rc;
    GridArgs a;
    zero_args(a);
    const int f = ctx->desc.n_features;
    const size_t esz = dtype == WX_F32 ? 4 : 8;
    for (int j = 0; j < f; ++j) {
        WX_CHECK_ARG(cols[j] != nullptr && (reinterpret_cast<uintptr_t>(cols[j]) % esz) == 0,
                     "wx_train_step_raw: column %d is NULL or misaligned", j);
        a.cols.col[j] = cols[j];
    }
    a.params = params;
    a.loader = dtype == WX_F32 ? LOAD_RAW_F32 : LOAD_RAW_F64;
    a.stats = stats;
    a.impute = impute_mode;
    a.y = y;
    a.row0 = row0;
    a.batch = batch;
    a.dropout_mode = dropout_mode;
    a.keep_mask = keep_mask;
    a.seed = seed;
    a.scale = 1.0f / float(batch);
    a.tail = TAIL_ADAM;
    a.p_rw = params;
    a.m = m;
    a.v = v;
    a.state = state;
    a.loss_out = loss_out;
    return launch_grid(ctx, a, MODE_TRAIN, (cudaStream_t)stream, "wx_train_step_raw");
}

int wx_train_steps_small(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state, const float* x,
                         const int64_t* y, const int64_t* row_idx, int64_t row0, int64_t batch, int64_t n_steps,
                         int dropout_mode, const uint64_t* keep_mask, uint64_t seed, float* loss_out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_train_steps_small: ctx is NULL");
    WX_CHECK_ARG(params && m && v && state && x && y && loss_out, "wx_train_steps_small: NULL pointer");
    WX_CHECK_ARG(batch >= 1 && batch <= WX_RESIDENT_MAX_BATCH,
                 "wx_train_steps_small: batch=%lld outside 1..%d (use wx_train_step)", (long long)batch,
                 WX_RESIDENT_MAX_BATCH);
    WX_CHECK_ARG(n_steps >= 0, "wx_train_steps_small: n_steps < 0");
    WX_CHECK_ARG(row_idx != nullptr || row0 >= 0, "wx_train_steps_small: row0 < 0");
    int rc = check_dropout(ctx, dropout_mode, keep_mask, "wx_train_steps_small");
    if (rc) return rc;
    if (n_steps == 0) return WX_OK;
    SmallArgs sa;
    memset(&sa, 0, sizeof(sa));
    sa.params = params;
    sa.m = m;
    sa.v = v;
    sa.state = state;
    sa.desc = ctx->desc;
    sa.x = x;
    sa.y = y;
    sa.row_idx = row_idx;
    sa.row0 = row0;
    sa.batch = batch;
    sa.n_steps = n_steps;
    sa.dropout_mode = dropout_mode;
    sa.keep_mask = keep_mask;
    sa.seed = seed;
    sa.loss_out = loss_out;
    int smem = 0;
    WX_CUDA(cudaSetDevice(ctx->device));
    if (ctx->peer.world > 1) {
        // after wx_peer_connect the steps are DDP steps: every rank launches the same number of steps of the same batch
        WX_CHECK_ARG(batch <= WARPS * R, "wx_train_steps_small: with connected peers the batch per rank is at most %d rows "
                     "(got %lld); larger batches take wx_tc_train_step", WARPS * R, (long long)batch);
        sa.peer = ctx->peer;
        sa.err = reinterpret_cast<int*>(ctx->scratch) + 8;
    }
    if (batch > WARPS * R) {
        // more rows than one CTA's eight warps take in a pass: the resident multi-CTA kernel (grid barrier per step)
        small_kernel_t kr = pick_resident_kernel(ctx->desc.n_features, &smem);
        WX_CHECK_ARG(kr != nullptr, "wx_train_steps_small: no kernel for F=%d", ctx->desc.n_features);
        const int64_t nchunks = (batch + R - 1) / R;
        int64_t g = (nchunks + WARPS - 1) / WARPS;                 // one chunk per warp while the CTAs last ...
        const int64_t gmax = ctx->num_sms < 32 ? ctx->num_sms : 32;  // ... but every CTA reads every slab: keep it to 32
        if (g > gmax) g = gmax;
        const int64_t cpc = (nchunks + g - 1) / g;
        g = (nchunks + cpc - 1) / cpc;                             // no CTA without work
        const size_t slab = (size_t(64) * ctx->desc.n_features + 194 + 1 + 3) & ~size_t(3);
        if (WX_SCRATCH_PART_OFF + 2 * size_t(g) * slab * 4 > ctx->scratch_bytes) {
            wx_set_error("wx_train_steps_small: scratch too small");
            return WX_ENOMEM;
        }
        sa.part = reinterpret_cast<float*>(ctx->scratch + WX_SCRATCH_PART_OFF);
        sa.bar = reinterpret_cast<unsigned long long*>(ctx->scratch + 128);     // bytes [128,136) of the ticket block
        sa.err = reinterpret_cast<int*>(ctx->scratch) + 8;
        cudaStream_t st = (cudaStream_t)stream;
        WX_CUDA(cudaMemsetAsync(sa.bar, 0, sizeof(unsigned long long), st));
        WX_CUDA(cudaFuncSetAttribute(reinterpret_cast<const void*>(kr), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        void* rargs[] = {&sa};
        WX_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(kr), dim3((unsigned)g), dim3(THREADS), rargs, smem, st));
        WX_LAUNCH_CHECK(ctx);
        return WX_OK;
    }
    small_kernel_t k = pick_small_kernel(ctx->desc.n_features, &smem);
    WX_CHECK_ARG(k != nullptr, "wx_train_steps_small: no kernel for F=%d", ctx->desc.n_features);
    WX_CUDA(cudaFuncSetAttribute(reinterpret_cast<const void*>(k), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    void* args[] = {&sa};
    WX_CUDA(cudaLaunchKernel(reinterpret_cast<const void*>(k), dim3(1), dim3(THREADS), args, smem, (cudaStream_t)stream));
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

int wx_mlp_eval(wx_ctx* ctx, const float* params, const float* x, const int64_t* y, const int64_t* row_idx,
                int64_t row0, int64_t batch, double* out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_mlp_eval: ctx is NULL");
    WX_CHECK_ARG(params && x && y && out, "wx_mlp_eval: NULL pointer");
    WX_CHECK_ARG(batch >= 0, "wx_mlp_eval: batch < 0");
    WX_CHECK_ARG(row_idx != nullptr || row0 >= 0, "wx_mlp_eval: row0 < 0");
    if (batch == 0) {
        WX_CUDA(cudaSetDevice(ctx->device));
        WX_CUDA(cudaMemsetAsync(out, 0, 2 * sizeof(double), (cudaStream_t)stream));
        return WX_OK;
    }
    GridArgs a;
    zero_args(a);
    a.params = params;
    a.loader = LOAD_ROWMAJOR;
    a.x = x;
    a.y = y;
    a.row_idx = row_idx;
    a.row0 = row0;
    a.batch = batch;
    a.dropout_mode = WX_DROPOUT_OFF;
    a.scale = 1.0f;
    a.eval_out = out;
    return launch_grid(ctx, a, MODE_EVAL, (cudaStream_t)stream, "wx_mlp_eval");
}

static int predict_common(wx_ctx* ctx, const float* params, const float* x, int64_t n, float* probs_out, void* stream,
                          bool logits, const char* who) {
    WX_CHECK_ARG(ctx != nullptr, "%s: ctx is NULL", who);
    WX_CHECK_ARG(n >= 0, "%s: n < 0", who);
    if (n == 0) return WX_OK;
    WX_CHECK_ARG(params && x && probs_out, "%s: NULL pointer", who);
    GridArgs a;
    zero_args(a);
    a.params = params;
    a.loader = LOAD_ROWMAJOR;
    a.x = x;
    a.y = nullptr;
    a.row0 = 0;
    a.batch = n;
    a.dropout_mode = WX_DROPOUT_OFF;
    a.scale = logits ? 0.0f : 1.0f;     // MODE_PREDICT reads `scale` as the output selector (process_chunk)
    a.probs_out = probs_out;
    return launch_grid(ctx, a, MODE_PREDICT, (cudaStream_t)stream, who);
}

int wx_mlp_predict(wx_ctx* ctx, const float* params, const float* x, int64_t n, float* probs_out, void* stream) {
    return predict_common(ctx, params, x, n, probs_out, stream, false, "wx_mlp_predict");
}

int wx_mlp_logits(wx_ctx* ctx, const float* params, const float* x, int64_t n, float* logits_out, void* stream) {
    return predict_common(ctx, params, x, n, logits_out, stream, true, "wx_mlp_logits");
}

}  // extern "C"

// =================================================================================================
// Allocation-free gradient entry point
// =================================================================================================
// Who this is for: hosts that may not allocate device memory at the moment they need a gradient --
// a plugin running under a framework that owns the whole GPU memory budget (a raw cudaMalloc there
// fails or defeats the framework's accounting), a real-time loop that cannot afford the implicit
// device synchronisation of cudaMalloc/cudaFree, or an embedding process whose policy forbids
// allocations after start-up.  wx_create() allocates (that is its job), so such callers cannot make a
// context; this entry point needs none.  Its scratch (the per-CTA partial-sum slabs and the arrival
// ticket that wx_mlp_grad keeps in the context) lives in the fixed array below, which is part of the
// library image: every GPU that loads libwxddp.so carries these 1.5 MiB until the process exits.
namespace {

constexpr size_t WX_STATIC_SCRATCH_BYTES = size_t(3) << 19;   // 1.5 MiB: ticket block + up to 148 slabs of the widest model
constexpr int64_t WX_STATIC_MAX_BATCH = int64_t(1) << 24;     // rows per call
constexpr int WX_STATIC_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char g_static_scratch[WX_STATIC_SCRATCH_BYTES];

struct StaticSlot {
    unsigned char* ptr = nullptr;   // DEVICE address of g_static_scratch on this device
    int num_sms = 0;
};
StaticSlot g_static_slot[WX_STATIC_MAX_DEVICES];
std::mutex g_static_mu;

// One-time set-up for the current device.  `g_static_scratch` as written in host code is only a
// host-side handle for the symbol; the kernels need the device address, which the runtime hands out
// per device through cudaGetSymbolAddress while that device is current.
int static_slot_for_current_device(StaticSlot** out, int* device_out) {
    int dev = -1;
    WX_CUDA(cudaGetDevice(&dev));
    WX_CHECK_ARG(dev >= 0 && dev < WX_STATIC_MAX_DEVICES, "wx_mlp_grad_static: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_static_mu);
    StaticSlot& s = g_static_slot[dev];
    if (s.ptr == nullptr) {
        cudaDeviceProp prop;
        WX_CUDA(cudaGetDeviceProperties(&prop, dev));
        if (prop.major != 10) {
            wx_set_error("wx_mlp_grad_static: device %d is sm_%d%d; this library is built for sm_100a only", dev,
                         prop.major, prop.minor);
            return WX_EARCH;
        }
        void* p = nullptr;
        WX_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        WX_CHECK_ARG(p != nullptr, "wx_mlp_grad_static: cudaGetSymbolAddress returned a null address");
        s.num_sms = prop.multiProcessorCount;
        s.ptr = static_cast<unsigned char*>(p);
    }
    *out = &s;
    *device_out = dev;
    return WX_OK;
}

}  // namespace

extern "C" int wx_mlp_grad_static(int n_features, const float* params, const float* x, const int64_t* y,
                                  const int64_t* row_idx, int64_t row0, int64_t batch, int dropout_mode,
                                  const uint64_t* keep_mask, uint64_t seed, uint64_t step, float scale, float* out,
                                  void* stream) {
    WX_CHECK_ARG(n_features >= 1 && n_features <= WX_MAX_FEATURES, "wx_mlp_grad_static: n_features=%d outside 1..%d",
                 n_features, WX_MAX_FEATURES);
    WX_CHECK_ARG(batch >= 1, "wx_mlp_grad_static: batch=%lld must be >= 1", (long long)batch);
    WX_CHECK_ARG(batch <= WX_STATIC_MAX_BATCH,
                 "wx_mlp_grad_static: batch=%lld exceeds the %lld-row limit of the fixed scratch entry point; split the "
                 "batch and add the outputs, or use wx_mlp_grad with a context",
                 (long long)batch, (long long)WX_STATIC_MAX_BATCH);
    WX_CHECK_ARG(params && x && y && out, "wx_mlp_grad_static: NULL params/x/y/out");
    WX_CHECK_ARG(row_idx != nullptr || row0 >= 0, "wx_mlp_grad_static: row0 < 0");
    WX_CHECK_ARG(dropout_mode == WX_DROPOUT_OFF || dropout_mode == WX_DROPOUT_MASK || dropout_mode == WX_DROPOUT_PHILOX,
                 "wx_mlp_grad_static: dropout_mode=%d", dropout_mode);
    WX_CHECK_ARG(dropout_mode != WX_DROPOUT_MASK || keep_mask != nullptr, "wx_mlp_grad_static: WX_DROPOUT_MASK needs keep_mask");
    StaticSlot* slot = nullptr;
    int dev = -1;
    int rc = static_slot_for_current_device(&slot, &dev);
    if (rc != WX_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // The array is shared by every call and still holds whatever the previous call left.  The slabs
    // are fully rewritten by each launch; the arrival ticket is the one thing that has to start at
    // zero, so it is cleared here (a fill, not an allocation) in case an earlier launch was aborted.
    WX_CUDA(cudaMemsetAsync(slot->ptr, 0, WX_SCRATCH_PART_OFF, st));
    // A context-shaped view of the static array on the stack, so the same launch path as wx_mlp_grad is used.
    wx_ctx view{};
    view.device = dev;
    view.num_sms = slot->num_sms;
    view.desc.n_features = n_features;
    view.desc.hidden = WX_HIDDEN;
    view.desc.classes = WX_CLASSES;
    view.desc.dropout_p = 0.2f;
    view.desc.lr = 0.01;
    view.desc.beta1 = 0.9;
    view.desc.beta2 = 0.999;
    view.desc.eps = 1e-8;
    view.n_params = int64_t(WX_HIDDEN) * n_features + WX_HIDDEN + WX_CLASSES * WX_HIDDEN + WX_CLASSES;
    view.scratch = slot->ptr;
    view.scratch_bytes = WX_STATIC_SCRATCH_BYTES;
    view.launches = 0;
    view.debug_times = nullptr;
    GridArgs a;
    zero_args(a);
    a.params = params;
    a.loader = LOAD_ROWMAJOR;
    a.x = x;
    a.y = y;
    a.row_idx = row_idx;
    a.row0 = row0;
    a.batch = batch;
    a.dropout_mode = dropout_mode;
    a.keep_mask = keep_mask;
    a.seed = seed;
    a.step = step;
    a.scale = scale;
    a.tail = TAIL_GRADS;
    a.out = out;
    return launch_grid(&view, a, MODE_TRAIN, st, "wx_mlp_grad_static");
}
