// One-shot all-reduce of the [P+1] gradient/loss vector over NVLink peer memory, executed INSIDE the step kernel's
// tail (replaces DDP's bucket all-reduce + the sync_dist scalar, train_lightning_ddp.py:70,:136).
//
// Every rank owns an "inbox" (cudaMalloc'ed, exported with CUDA IPC, mapped by all peers):
//     [ flags: uint64 [2 parities][8 ranks] | data: float [2 parities][8 ranks][slab stride] ]
// Step t (tag t, parity t & 1): the last CTA of rank r stores its local sums into slot [parity][r] of EVERY rank's
// inbox (plain stores to mapped peer addresses -> NVLink / NVSwitch), fences system-wide, publishes tag t in each
// inbox's flag [parity][r], then waits until all W flags of its own inbox carry tag t and adds the W vectors in rank
// order -- the same order on every rank, so the replicas stay bit-identical.  Two parities are enough: a rank cannot
// get more than one step ahead of a peer because it needs that peer's flag for the current step first.
// The wait is bounded; on expiry the caller records an error code instead of hanging the GPU.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace wxpeer {

constexpr int MAXW = 8;
constexpr int FLAG_BYTES = 256;                 // 2 x 8 x uint64 = 128 B, padded
constexpr int SLOT_FLOATS = 2304;               // >= largest slab stride (64*32+195 -> 2244), multiple of 4
constexpr size_t INBOX_BYTES = FLAG_BYTES + size_t(2) * MAXW * SLOT_FLOATS * 4;

struct PeerArgs {
    int world;                                  // 0/1 = no exchange
    int rank;
    unsigned char* inbox[MAXW];                 // inbox of rank q as mapped in THIS process (inbox[rank] = own)
};

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_volatile_f32(const float* p) {
    float v;
    asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

// vec: n floats in SHARED memory holding this rank's sums on entry, the all-rank sums on exit.  Called by all `nthreads`
// threads of one CTA (barrier id 0).  Returns false on timeout.
__device__ __forceinline__ bool allreduce_cta(float* vec, int n, const PeerArgs& pa, uint64_t tag, int tid, int nthreads) {
    const int par = int(tag & 1);
    const int w = pa.world;
    // 1. push my vector into slot [par][rank] of every inbox (own included)
    for (int q = 0; q < w; ++q) {
        float* dst = reinterpret_cast<float*>(pa.inbox[q] + FLAG_BYTES) + (size_t(par) * MAXW + pa.rank) * SLOT_FLOATS;
        for (int i = tid; i < n; i += nthreads) dst[i] = vec[i];
    }
    __threadfence_system();
    __syncthreads();
    // 2. publish
    if (tid < w) st_release_sys(reinterpret_cast<uint64_t*>(pa.inbox[tid]) + par * MAXW + pa.rank, tag);
    // 3. wait for everybody's slot in my own inbox
    __shared__ int ok_flag;
    if (tid == 0) ok_flag = 1;
    __syncthreads();
    if (tid < w) {
        const uint64_t* f = reinterpret_cast<const uint64_t*>(pa.inbox[pa.rank]) + par * MAXW + tid;
        bool got = false;
        for (long long spin = 0; spin < (1ll << 24); ++spin) {
            if (ld_acquire_sys(f) == tag) {
                got = true;
                break;
            }
            __nanosleep(100);
        }
        if (!got) ok_flag = 0;
    }
    __syncthreads();
    // 4. sum in rank order
    const float* base = reinterpret_cast<const float*>(pa.inbox[pa.rank] + FLAG_BYTES) + size_t(par) * MAXW * SLOT_FLOATS;
    for (int i = tid; i < n; i += nthreads) {
        float s = 0.0f;
        for (int q = 0; q < w; ++q) s += ld_volatile_f32(base + size_t(q) * SLOT_FLOATS + i);
        vec[i] = s;
    }
    __syncthreads();
    return ok_flag != 0;
}

}  // namespace wxpeer
