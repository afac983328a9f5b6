// Context, error string and library identity for libwxddp.so.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"
#include "wxddp_dev.h"

static thread_local char g_err[512] = "";

void wx_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" {

int wx_version(void) { return 100; }   // 0.1.0
const char* wx_build_arch(void) { return "sm_100a"; }
const char* wx_last_error(void) { return g_err; }

int wx_create(int device, const wx_model_desc* desc, wx_ctx** out) {
    WX_CHECK_ARG(out != nullptr && desc != nullptr, "wx_create: null argument");
    *out = nullptr;
    WX_CHECK_ARG(desc->n_features >= 1 && desc->n_features <= WX_MAX_FEATURES,
                 "wx_create: n_features=%d outside 1..%d", desc->n_features, WX_MAX_FEATURES);
    WX_CHECK_ARG(desc->hidden == WX_HIDDEN, "wx_create: hidden=%d, only %d is built (train_lightning_ddp.py:58)",
                 desc->hidden, WX_HIDDEN);
    WX_CHECK_ARG(desc->classes == WX_CLASSES, "wx_create: classes=%d, only %d is built (train_lightning_ddp.py:61)",
                 desc->classes, WX_CLASSES);
    WX_CHECK_ARG(desc->dropout_p == 0.0f || desc->dropout_p == 0.2f,
                 "wx_create: dropout_p=%g, only 0.2 (train_lightning_ddp.py:60) and 0 are built", desc->dropout_p);
    WX_CHECK_ARG(desc->optimizer == WX_OPT_ADAM || desc->optimizer == WX_OPT_SGD, "wx_create: optimizer=%d is neither WX_OPT_ADAM "
                 "nor WX_OPT_SGD", desc->optimizer);
    WX_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    WX_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        wx_set_error("wx_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major,
                     prop.minor);
        return WX_EARCH;
    }
    wx_ctx* c = new wx_ctx();
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    c->desc = *desc;
    c->n_params = int64_t(WX_HIDDEN) * desc->n_features + WX_HIDDEN + WX_CLASSES * WX_HIDDEN + WX_CLASSES;
    c->scratch_bytes = WX_SCRATCH_BYTES;
    c->launches = 0;
    c->wide_reserved_sms = 0;
    c->debug_times = nullptr;
    c->inbox = nullptr;
    c->peer_epoch = 0;
    memset(&c->peer, 0, sizeof(c->peer));
    memset(c->peer_mapped, 0, sizeof(c->peer_mapped));
    cudaError_t e = cudaMalloc(&c->scratch, c->scratch_bytes);
    if (e != cudaSuccess) {
        wx_set_error("wx_create: cudaMalloc(%zu) failed: %s", c->scratch_bytes, cudaGetErrorString(e));
        delete c;
        return WX_ECUDA;
    }
    e = cudaMemset(c->scratch, 0, c->scratch_bytes);
    if (e != cudaSuccess) {
        wx_set_error("wx_create: cudaMemset failed: %s", cudaGetErrorString(e));
        cudaFree(c->scratch);
        delete c;
        return WX_ECUDA;
    }
    *out = c;
    return WX_OK;
}

int wx_destroy(wx_ctx* ctx) {
    if (!ctx) return WX_OK;
    cudaSetDevice(ctx->device);
    for (int q = 0; q < wxpeer::MAXW; ++q)
        if (ctx->peer_mapped[q]) cudaIpcCloseMemHandle(ctx->peer_mapped[q]);
    if (ctx->inbox) cudaFree(ctx->inbox);
    cudaFree(ctx->scratch);
    delete ctx;
    return WX_OK;
}

int64_t wx_n_params(const wx_ctx* ctx) { return ctx ? ctx->n_params : -1; }
int wx_num_sms(const wx_ctx* ctx) { return ctx ? ctx->num_sms : -1; }
int64_t wx_launch_count(const wx_ctx* ctx) { return ctx ? ctx->launches : -1; }

int wx_peer_export(wx_ctx* ctx, void* handle_out) {
    WX_CHECK_ARG(ctx != nullptr && handle_out != nullptr, "wx_peer_export: NULL argument");
    WX_CUDA(cudaSetDevice(ctx->device));
    if (ctx->inbox == nullptr) {
        WX_CUDA(cudaMalloc(&ctx->inbox, wxpeer::INBOX_BYTES));
        WX_CUDA(cudaMemset(ctx->inbox, 0, wxpeer::INBOX_BYTES));
        WX_CUDA(cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t h;
    WX_CUDA(cudaIpcGetMemHandle(&h, ctx->inbox));
    static_assert(sizeof(h) == 64, "CUDA IPC handle is 64 bytes");
    memcpy(handle_out, &h, sizeof(h));
    return WX_OK;
}

int wx_peer_connect(wx_ctx* ctx, int rank, int world, const void* handles) {
    WX_CHECK_ARG(ctx != nullptr && handles != nullptr, "wx_peer_connect: NULL argument");
    WX_CHECK_ARG(world >= 1 && world <= wxpeer::MAXW && rank >= 0 && rank < world, "wx_peer_connect: rank=%d world=%d (max %d)",
                 rank, world, wxpeer::MAXW);
    WX_CHECK_ARG(ctx->inbox != nullptr, "wx_peer_connect: call wx_peer_export first");
    WX_CUDA(cudaSetDevice(ctx->device));
    wxpeer::PeerArgs pa;
    memset(&pa, 0, sizeof(pa));
    pa.world = world;
    pa.rank = rank;
    for (int q = 0; q < world; ++q) {
        if (q == rank) {
            pa.inbox[q] = ctx->inbox;
            continue;
        }
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const unsigned char*>(handles) + size_t(q) * 64, 64);
        void* p = nullptr;
        WX_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        ctx->peer_mapped[q] = p;
        pa.inbox[q] = static_cast<unsigned char*>(p);
    }
    pa.epoch = ++ctx->peer_epoch;            // every rank connects equally often: the epochs agree across ranks
    if (pa.epoch == 0 || pa.epoch > 255) pa.epoch = ctx->peer_epoch = 1;
    ctx->peer = pa;
    return WX_OK;
}

int wxdev_cta_times(wx_ctx* ctx, long long* device_buf) {
    if (!ctx) return WX_EINVAL;
    ctx->debug_times = device_buf;
    return WX_OK;
}

}  // extern "C"
