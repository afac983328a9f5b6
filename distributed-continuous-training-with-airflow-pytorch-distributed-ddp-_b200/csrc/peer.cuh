// One-shot all-reduce of the [P+1] gradient/loss vector over NVLink peer memory, executed INSIDE the step kernel's
// tail (replaces DDP's bucket all-reduce + the sync_dist scalar, train_lightning_ddp.py:70,:136).
//
// Every rank owns an "inbox" (cudaMalloc'ed, exported with CUDA IPC, mapped by all peers):
//     uint64 [2 parities][8 ranks][SLOT_FLOATS]      one 8-byte word per element: { fp32 value | 32-bit tag }
// Step t (parity t & 1): the last CTA of rank r stores its local sums, each packed with the step's tag into ONE 8-byte
// word (single-copy atomic), into slot [parity][r] of EVERY rank's inbox -- plain stores to mapped peer addresses, no
// fence, no separate flag: a word whose tag matches carries its value.  It then polls the W slots of its own inbox
// element by element and adds the W values in rank order -- the same order on every rank, so the replicas stay
// bit-identical.  Two parities are enough: a rank cannot get more than one step ahead of a peer because it needs that
// peer's words for the current step first.  tag = (connection epoch << 24) | (step & 0xffffff), never 0: words of an
// earlier step, an earlier connection or a fresh (zeroed) inbox never match.
// The wait is bounded; on expiry the caller skips the update and records an error code instead of hanging the GPU.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace wxpeer {

constexpr int MAXW = 8;
constexpr int SLOT_FLOATS = 2304;               // >= largest slab stride (64*32+195 -> 2244), multiple of 4
constexpr size_t INBOX_BYTES = size_t(2) * MAXW * SLOT_FLOATS * 8;

struct PeerArgs {
    int world;                                  // 0/1 = no exchange
    int rank;
    unsigned int epoch;                         // connection count of this context (1, 2, ...): top byte of every tag
    unsigned char* inbox[MAXW];                 // inbox of rank q as mapped in THIS process (inbox[rank] = own)
};

__device__ __forceinline__ void st_relaxed_sys_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_sys_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// vec: n floats in SHARED memory holding this rank's sums on entry, the all-rank sums on exit.  Called by all `nthreads`
// threads of one CTA (barrier id 0).  `step` is the 1-based step being taken.  Returns false on timeout (vec is then
// unspecified and the caller must not apply the step).
__device__ __forceinline__ bool allreduce_cta(float* vec, int n, const PeerArgs& pa, uint64_t step, int tid, int nthreads) {
    const int par = int(step & 1);
    const int w = pa.world;
    const uint32_t tag = (pa.epoch << 24) | uint32_t(step & 0xffffffu);
    // 1. push my vector, value and tag in one word, into slot [par][rank] of every inbox (own included)
    for (int i = tid; i < n; i += nthreads) {
        const uint64_t word = (uint64_t(tag) << 32) | uint64_t(__float_as_uint(vec[i]));
        for (int q = 0; q < w; ++q) {
            uint64_t* dst = reinterpret_cast<uint64_t*>(pa.inbox[q]) + (size_t(par) * MAXW + pa.rank) * SLOT_FLOATS + i;
            st_relaxed_sys_u64(dst, word);
        }
    }
    // 2. collect: every element of every rank's slot in my own inbox, added in rank order
    __shared__ int ok_flag;
    if (tid == 0) ok_flag = 1;
    __syncthreads();
    const uint64_t* base = reinterpret_cast<const uint64_t*>(pa.inbox[pa.rank]) + size_t(par) * MAXW * SLOT_FLOATS;
    bool ok = true;
    for (int i = tid; i < n; i += nthreads) {
        // all W words of this element are requested at once (each load is a ~1 us round trip: one after the other they
        // cost 17 us at 8 ranks); a word whose tag is not there yet is then polled on its own
        uint64_t word[MAXW];
#pragma unroll
        for (int q = 0; q < MAXW; ++q) word[q] = q < w ? ld_relaxed_sys_u64(base + size_t(q) * SLOT_FLOATS + i) : 0ull;
        float s = 0.0f;
#pragma unroll
        for (int q = 0; q < MAXW; ++q) {
            if (q < w) {
                const uint64_t* src = base + size_t(q) * SLOT_FLOATS + i;
                uint64_t wd = word[q];
                for (long long spin = 0; uint32_t(wd >> 32) != tag && spin < (1ll << 22); ++spin) {
                    __nanosleep(64);
                    wd = ld_relaxed_sys_u64(src);
                }
                if (uint32_t(wd >> 32) != tag) ok = false;
                s += __uint_as_float(uint32_t(wd));
            }
        }
        vec[i] = s;
    }
    if (!ok) ok_flag = 0;
    __syncthreads();
    return ok_flag != 0;
}

}  // namespace wxpeer
