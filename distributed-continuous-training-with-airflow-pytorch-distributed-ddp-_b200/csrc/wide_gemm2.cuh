// Two-SM variant of the wide bf16 GEMM (wide_gemm.cuh): clusters of 2 CTAs on one TPC compute a 256 x 256 output tile with
// tcgen05.mma.cta_group::2.  Each CTA stages ITS 128 rows of A and ITS 128-column half of B (32 KiB per stage instead of
// 48), the leader CTA's MMA warp issues one M256 N256 K16 instruction that reads both CTAs' shared memory, and each CTA
// keeps the accumulator of its own 128 rows (all 256 columns) in its own TMEM.  Per FLOP the pair reads a third less
// shared memory than two independent 128 x 256 tiles, and the ring is 6 stages deep.
//
// Barrier protocol (per stage / accumulator buffer):
//   full[s]        lives in the LEADER: both CTAs' TMA loads complete_tx on it (barrier address with the peer bit cleared);
//                  only the leader's producers arrive.expect_tx (for both CTAs' bytes)
//   empty[s]       one per CTA: the leader's tcgen05.commit multicasts the arrive to both
//   tmem_full[a]   one per CTA, multicast commit as well
//   tmem_empty[a]  in the leader, 2 x 8 arrivals: the peer's epilogue warps arrive remotely (mapa + mbarrier.arrive.cluster)
// Used for the big products of the wide MLP (M % 256 == 0, N % 256 == 0, no split-K); everything else takes the 1-CTA kernel.
#pragma once

#include "wide_gemm.cuh"

namespace wxwide {

constexpr int STAGES2 = 6;
constexpr int B2_STAGE_BYTES = (BN / 2) * BK * 2;   // 16 KiB: this CTA's half of the B tile
constexpr size_t GEMM2_SMEM = size_t(STAGES2) * (A_STAGE_BYTES + B2_STAGE_BYTES) + 1024;
constexpr uint32_t PEER_BIT_MASK = 0xFEFFFFFFu;     // shared::cluster address of the same offset in CTA 0 of the pair

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma2_load_2d(void* smem_dst, const CUtensorMap* tm, int32_t c0, int32_t c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(wxtc::smem_u32(smem_dst)), "l"(tm), "r"(c0), "r"(c1), "r"(wxtc::smem_u32(bar) & PEER_BIT_MASK)
        : "memory");
}
__device__ __forceinline__ void tma2_load_3d(void* smem_dst, const CUtensorMap* tm, int32_t c0, int32_t c1, int32_t c2,
                                             uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(wxtc::smem_u32(smem_dst)), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(wxtc::smem_u32(bar) & PEER_BIT_MASK)
        : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    const uint32_t acc = accumulate ? 1u : 0u;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs of the pair once all MMAs issued so far have completed
__device__ __forceinline__ void umma2_commit_both(uint64_t* bar) {
    const uint16_t mask = 3;
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(wxtc::smem_u32(bar)), "h"(mask)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(wxtc::smem_u32(bar)), "r"(0));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ void tmem2_alloc(uint32_t* smem_out, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(wxtc::smem_u32(smem_out)), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem2_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

template <bool A_K, bool B_K, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
wide_gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmArgs g) {
    extern __shared__ unsigned char wide_smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(wide_smem_raw) + 1023) & ~uintptr_t(1023));
    unsigned char* sA = smem;
    unsigned char* sB = smem + size_t(STAGES2) * A_STAGE_BYTES;
    __shared__ uint64_t full[STAGES2], empty[STAGES2], tmem_full[2], tmem_empty[2];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();                 // 0 = leader
    const int64_t tiles_m = g.M / (2 * BM), tiles_n = g.N / BN;
    const int64_t mn_tiles = tiles_m * tiles_n;
    const int64_t n_tiles = mn_tiles * g.split_k;            // split-K: work unit = (split, tile), one output slab per split
    const int64_t cl = blockIdx.x >> 1, n_cl = gridDim.x >> 1;
    const int kb_total = int(g.K / BK);
    const int kb_per = (kb_total + g.split_k - 1) / g.split_k;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2; ++s) {
            wxtc::mbar_init(&full[s], 2);                    // the leader's two producers
            wxtc::mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            wxtc::mbar_init(&tmem_full[a], 1);
            wxtc::mbar_init(&tmem_empty[a], 2 * EPI_WARPS);
        }
        wxtc::fence_mbar_init();
    }
    if (warp == 2) tmem2_alloc(&tmem_base_s, 512);
    wxtc::tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                       // both CTAs' barriers exist before anyone signals them
    wxtc::tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    bool ok = true;

    if (warp < 2) {
        // ================= TMA producers (both CTAs): warp 0 loads this CTA's A rows, warp 1 its half of B =================
        if (lane == 0) {
            const bool is_a = warp == 0;
            uint32_t it = 0;
            for (int64_t tile = cl; tile < n_tiles && ok; tile += n_cl) {
                const int64_t mn = tile % mn_tiles;
                const int split = int(tile / mn_tiles);
                const int32_t m0 = int32_t(mn / tiles_n) * 2 * BM + int32_t(rank) * BM;
                const int32_t n0 = int32_t(mn % tiles_n) * BN + int32_t(rank) * (BN / 2);
                const int kb0 = split * kb_per, kb1 = (kb0 + kb_per < kb_total) ? kb0 + kb_per : kb_total;
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int s = it % STAGES2;
                    if (it >= STAGES2 && !wxtc::mbar_wait_spin(&empty[s], ((it / STAGES2) - 1) & 1)) { ok = false; break; }
                    if (is_a) {
                        if (rank == 0) wxtc::mbar_arrive_expect_tx(&full[s], 2 * A_STAGE_BYTES);
                        unsigned char* a = sA + size_t(s) * A_STAGE_BYTES;
                        if (A_K) tma2_load_2d(a, &tmA, kb * BK, m0, &full[s]);
                        else tma2_load_3d(a, &tmA, 0, kb * BK, m0 / 64, &full[s]);
                    } else {
                        if (rank == 0) wxtc::mbar_arrive_expect_tx(&full[s], 2 * B2_STAGE_BYTES);
                        unsigned char* b = sB + size_t(s) * B2_STAGE_BYTES;
                        if (B_K) tma2_load_2d(b, &tmB, kb * BK, n0, &full[s]);
                        else tma2_load_3d(b, &tmB, 0, kb * BK, n0 / 64, &full[s]);
                    }
                }
            }
            if (!ok) atomicExch(g.err, 211);
        }
    } else if (warp == 2) {
        // ================= MMA issuer: the leader CTA only =================
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, !A_K, !B_K);
            uint32_t it = 0, local = 0;
            for (int64_t tile = cl; tile < n_tiles && ok; tile += n_cl, ++local) {
                const uint32_t acc = local & 1;
                if (local >= 2 && !wxtc::mbar_wait_spin(&tmem_empty[acc], ((local >> 1) - 1) & 1)) { ok = false; break; }
                wxtc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                const int split = int(tile / mn_tiles);
                const int kb0 = split * kb_per, kb1 = (kb0 + kb_per < kb_total) ? kb0 + kb_per : kb_total;
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int s = it % STAGES2;
                    if (!wxtc::mbar_wait_spin(&full[s], (it / STAGES2) & 1)) { ok = false; break; }
                    wxtc::tc_fence_after();
                    const uint32_t a_addr = wxtc::smem_u32(sA + size_t(s) * A_STAGE_BYTES);
                    const uint32_t b_addr = wxtc::smem_u32(sB + size_t(s) * B2_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        const uint64_t ad = A_K ? wxtc::make_smem_desc(a_addr + k * 32, 16, 1024)
                                                : wxtc::make_smem_desc(a_addr + k * 2048, 8192, 1024);
                        const uint64_t bd = B_K ? wxtc::make_smem_desc(b_addr + k * 32, 16, 1024)
                                                : wxtc::make_smem_desc(b_addr + k * 2048, 8192, 1024);
                        umma2_bf16(d_tmem, ad, bd, idesc, ((kb - kb0) | k) != 0);
                    }
                    umma2_commit_both(&empty[s]);
                }
                if (!ok) break;
                umma2_commit_both(&tmem_full[acc]);
            }
            if (!ok) atomicExch(g.err, 212);
        }
    } else {
        // ================= epilogue warps (both CTAs): this CTA's 128 rows x 256 columns =================
        const int q = warp & 3;
        const int half = (warp - 3) >> 2;
        uint32_t local = 0;
        for (int64_t tile = cl; tile < n_tiles; tile += n_cl, ++local) {
            const uint32_t acc = local & 1;
            const int64_t mn = tile % mn_tiles;
            const int64_t m0 = (mn / tiles_n) * 2 * BM + int64_t(rank) * BM, n0 = (mn % tiles_n) * BN;
            const int64_t out_off = (tile / mn_tiles) * g.split_stride;
            if (!wxtc::mbar_wait(&tmem_full[acc], (local >> 1) & 1)) {
                if (lane == 0) atomicExch(g.err, 213);
                break;
            }
            wxtc::tc_fence_after();
            const int64_t row = m0 + q * 32 + lane;
            const bool row_ok = row < g.M;
#pragma unroll 1
            for (int c = half * (BN / 32); c < (half + 1) * (BN / 32); ++c) {
                float v[16];
                wxtc::tmem_ld16(tmem_base + (uint32_t(q * 32) << 16) + acc * BN + c * 16, v);
                wxtc::tmem_ld_wait();
                epilogue_chunk<EPI>(v, g, row, row_ok, n0 + c * 16, out_off);
                __syncwarp();
            }
            wxtc::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (rank == 0) wxtc::mbar_arrive(&tmem_empty[acc]);
                else mbar_arrive_leader(&tmem_empty[acc]);
            }
        }
    }
    wxtc::tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                       // nobody leaves while the pair's MMAs / remote arrives are in flight
    if (warp == 2) tmem2_dealloc(tmem_base, 512);
}

}  // namespace wxwide
