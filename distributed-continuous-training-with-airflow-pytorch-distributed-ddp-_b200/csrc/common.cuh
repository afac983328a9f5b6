// Shared declarations for libwxddp.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "wxddp.h"
#include "peer.cuh"

#define WX_WARP 32

// Scratch layout inside wx_ctx::scratch (bytes).  One cudaMalloc in wx_create, never resized.
//   [0, 256)                      tickets / counters (zeroed at create; every kernel that uses one resets it)
//   [WX_SCRATCH_PART_OFF, ...)    per-CTA partial slabs of the step kernels and the stats kernel
static constexpr size_t WX_SCRATCH_TICKET_BYTES = 256;
static constexpr size_t WX_SCRATCH_PART_OFF = 256;
static constexpr size_t WX_SCRATCH_BYTES = size_t(16) << 20;   // 16 MiB

static constexpr int WX_MAX_GRID_PARTS = 1184;                  // 148 SMs x 8: upper bound on partial slabs

struct wx_ctx {
    int device;
    int num_sms;
    wx_model_desc desc;
    int64_t n_params;
    unsigned char* scratch;      // device
    size_t scratch_bytes;
    int64_t launches;
    long long* debug_times;   // optional device buffer [grid][4]: smid, t_start, t_loop_end, t_exit (globaltimer ns)
    unsigned char* inbox;     // this rank's peer inbox (wx_peer_export), NULL until exported
    unsigned int peer_epoch;  // number of wx_peer_connect calls so far (part of every exchange tag)
    wxpeer::PeerArgs peer;    // world <= 1 until wx_peer_connect
    void* peer_mapped[wxpeer::MAXW];   // cudaIpcOpenMemHandle results to close at destroy
    int wide_reserved_sms;    // SMs the wide GEMMs leave free for a concurrent collective (wx_wide_reserve_sms), default 0
};

void wx_set_error(const char* fmt, ...);

#define WX_CHECK_ARG(cond, ...)            \
    do {                                   \
        if (!(cond)) {                     \
            wx_set_error(__VA_ARGS__);     \
            return WX_EINVAL;              \
        }                                  \
    } while (0)

#define WX_CUDA(call)                                                                     \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            wx_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return WX_ECUDA;                                                              \
        }                                                                                 \
    } while (0)

#define WX_LAUNCH_CHECK(ctx)                                                              \
    do {                                                                                  \
        cudaError_t e__ = cudaGetLastError();                                             \
        if (e__ != cudaSuccess) {                                                         \
            wx_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
            return WX_ECUDA;                                                              \
        }                                                                                 \
        (ctx)->launches += 1;                                                             \
    } while (0)

// pointers to up to 32 columns, passed by value as a kernel argument
struct WxCols {
    const void* col[WX_MAX_FEATURES];
    const uint8_t* valid[WX_MAX_FEATURES];
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
