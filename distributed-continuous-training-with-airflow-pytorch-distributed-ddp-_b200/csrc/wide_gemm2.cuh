// Two-SM variant of the wide bf16 GEMM (wide_gemm.cuh): clusters of 2 CTAs on one TPC compute a 256 x 256 (NB = 1) or
// 256 x 512 (NB = 2) output tile with tcgen05.mma.cta_group::2.  Each CTA stages ITS 128 rows of A and ITS 128-column half
// of every 256-column block of B, the leader CTA's MMA warp issues M256 N256 K16 instructions that read both CTAs' shared
// memory, and each CTA keeps the accumulator of its own 128 rows in its own TMEM.
//
// Barrier protocol (per stage / accumulator buffer):
//   full[s]        lives in the LEADER: both CTAs' TMA loads complete_tx on it (barrier address with the peer bit cleared);
//                  only the leader's producers arrive.expect_tx (for both CTAs' bytes)
//   empty[s]       one per CTA: the leader's tcgen05.commit multicasts the arrive to both
//   tmem_full[a]   one per CTA, multicast commit as well
//   tmem_empty[a]  in the leader, 2 x EW arrivals: the peer's epilogue warps arrive remotely (mapa + mbarrier.arrive.cluster)
// Used for the big products of the wide MLP (M % 256 == 0, N % 256 == 0); everything else takes the 1-CTA kernel.
//
// NB = 2 (two N = 256 MMAs per K step against the same A rows; 4 stages of 48 KiB).  A 256 x 256 tile pulls 64 KiB per
// 64-deep K block and pair, i.e. 64 B per clock and SM at the tensor pipe's full rate -- 9.5 KB per clock for the chip, more
// than the L2 delivers -- so the narrow tile (6 stages of 32 KiB, two accumulators) is L2-bound at roughly 2/3 of the
// tensor rate.  The wide tile needs 48 B per clock and SM.  It costs the second accumulator (512 TMEM columns hold one
// 128 x 512 fp32 tile), so the epilogue of a tile is no longer hidden behind the next tile's MMAs; the loads of the next
// tile still run ahead, and the epilogue warps do everything that does not need the accumulator before they wait for it.
//
// Work units: output tiles, optionally cut along K (GemmArgs::split_k: every tile; GemmArgs::tail_*: only the tiles of the
// last, partly filled round -- see decode() below).
#pragma once

#include "wide_gemm.cuh"

namespace wxwide {

constexpr int B2_STAGE_BYTES = (BN / 2) * BK * 2;   // 16 KiB: this CTA's half of one 256-column B block
constexpr int EPI_STAGE_BYTES = 32 * 128;            // per epilogue warp: 32 rows x 128 bytes of output, 16-byte units XOR-swizzled
// EW = epilogue warps per CTA: 8 for the long products; 16 (and a shallower ring) for products whose contraction is a
// K block or two, where the kernel is bound by the epilogue's instructions, not by the MMAs.
__host__ __device__ constexpr int gemm2_stages(int nb, int ew) { return ew > 8 ? 4 : (nb == 1 ? 6 : 4); }
__host__ __device__ constexpr size_t gemm2_smem(int nb, int ew) {
    return size_t(gemm2_stages(nb, ew)) * (A_STAGE_BYTES + nb * B2_STAGE_BYTES) + size_t(ew) * EPI_STAGE_BYTES + 1024;
}
__host__ __device__ constexpr int gemm2_threads(int ew) { return (3 + ew) * 32; }
__device__ __forceinline__ uint32_t epi_unit_addr(uint32_t base, int row, int unit) {
    return base + uint32_t(row) * 128u + (uint32_t(unit ^ (row & 7)) << 4);
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
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

template <bool A_K, bool B_K, int EPI, int NB, int EW>
__global__ void __launch_bounds__(gemm2_threads(EW), 1)
wide_gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmArgs g) {
    extern __shared__ unsigned char wide_smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(wide_smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int STAGES2 = gemm2_stages(NB, EW);
    constexpr int BNT = NB * BN;                              // tile width
    constexpr int BT_STAGE_BYTES = NB * B2_STAGE_BYTES;
    constexpr uint32_t NACC = 2 / NB;                         // accumulator buffers that fit in 512 TMEM columns
    unsigned char* sA = smem;
    unsigned char* sB = smem + size_t(STAGES2) * A_STAGE_BYTES;
    unsigned char* sEpi = sB + size_t(STAGES2) * BT_STAGE_BYTES;
    __shared__ uint64_t full[STAGES2], empty[STAGES2], tmem_full[2], tmem_empty[2];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();                 // 0 = leader
    const int64_t tiles_m = g.M / (2 * BM), tiles_n = g.N / BNT;
    const int64_t mn_tiles = tiles_m * tiles_n;
    const bool tail = g.tail_split > 1;
    // work units: (split, tile) with one output slab per split; or whole tiles followed by the K-ranges of the tail tiles
    const int64_t n_tiles = tail ? g.tail_first + (mn_tiles - g.tail_first) * g.tail_split : mn_tiles * g.split_k;
    const int64_t cl = blockIdx.x >> 1, n_cl = gridDim.x >> 1;
    const int kb_total = int(g.K / BK);
    const int n_ranges = tail ? g.tail_split : g.split_k;
    const int kb_per = (kb_total + n_ranges - 1) / n_ranges;
    auto decode = [&](int64_t unit, int64_t& mn, int& kb0, int& kb1, int64_t& out_off) {
        int range = 0;
        if (!tail) {
            mn = unit % mn_tiles;
            range = int(unit / mn_tiles);
            out_off = range * g.split_stride;
        } else if (unit < g.tail_first) {
            mn = unit;
            kb0 = 0;
            kb1 = kb_total;
            out_off = 0;
            return;
        } else {
            const int64_t r = unit - g.tail_first;
            mn = g.tail_first + r / g.tail_split;
            range = int(r % g.tail_split);
            out_off = range == 0 ? 0 : g.tail_off + (range - 1) * g.tail_stride;
        }
        kb0 = range * kb_per;
        kb1 = (kb0 + kb_per < kb_total) ? kb0 + kb_per : kb_total;
    };

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2; ++s) {
            wxtc::mbar_init(&full[s], 2);                    // the leader's two producers
            wxtc::mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            wxtc::mbar_init(&tmem_full[a], 1);
            wxtc::mbar_init(&tmem_empty[a], 2 * EW);
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
                int64_t mn, out_off;
                int kb0, kb1;
                decode(tile, mn, kb0, kb1, out_off);
                const int32_t m0 = int32_t(mn / tiles_n) * 2 * BM + int32_t(rank) * BM;
                const int32_t n0 = int32_t(mn % tiles_n) * BNT + int32_t(rank) * (BN / 2);
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int s = it % STAGES2;
                    if (it >= STAGES2 && !wxtc::mbar_wait_spin(&empty[s], ((it / STAGES2) - 1) & 1)) { ok = false; break; }
                    if (is_a) {
                        if (rank == 0) wxtc::mbar_arrive_expect_tx(&full[s], 2 * A_STAGE_BYTES);
                        unsigned char* a = sA + size_t(s) * A_STAGE_BYTES;
                        if (A_K) tma2_load_2d(a, &tmA, kb * BK, m0, &full[s]);
                        else tma2_load_3d(a, &tmA, 0, kb * BK, m0 / 64, &full[s]);
                    } else {
                        if (rank == 0) wxtc::mbar_arrive_expect_tx(&full[s], 2 * BT_STAGE_BYTES);
#pragma unroll
                        for (int j = 0; j < NB; ++j) {
                            unsigned char* b = sB + size_t(s) * BT_STAGE_BYTES + j * B2_STAGE_BYTES;
                            if (B_K) tma2_load_2d(b, &tmB, kb * BK, n0 + j * BN, &full[s]);
                            else tma2_load_3d(b, &tmB, 0, kb * BK, (n0 + j * BN) / 64, &full[s]);
                        }
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
                const uint32_t acc = local % NACC, use = local / NACC;
                if (use >= 1 && !wxtc::mbar_wait_spin(&tmem_empty[acc], (use - 1) & 1)) { ok = false; break; }
                wxtc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                int64_t mn, out_off;
                int kb0, kb1;
                decode(tile, mn, kb0, kb1, out_off);
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int s = it % STAGES2;
                    if (!wxtc::mbar_wait_spin(&full[s], (it / STAGES2) & 1)) { ok = false; break; }
                    wxtc::tc_fence_after();
                    const uint32_t a_addr = wxtc::smem_u32(sA + size_t(s) * A_STAGE_BYTES);
                    const uint32_t b_addr = wxtc::smem_u32(sB + size_t(s) * BT_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        const uint64_t ad = A_K ? wxtc::make_smem_desc(a_addr + k * 32, 16, 1024)
                                                : wxtc::make_smem_desc(a_addr + k * 2048, 8192, 1024);
#pragma unroll
                        for (int j = 0; j < NB; ++j) {
                            const uint32_t bj = b_addr + j * B2_STAGE_BYTES;
                            const uint64_t bd = B_K ? wxtc::make_smem_desc(bj + k * 32, 16, 1024)
                                                    : wxtc::make_smem_desc(bj + k * 2048, 8192, 1024);
                            umma2_bf16(d_tmem + j * BN, ad, bd, idesc, ((kb - kb0) | k) != 0);
                        }
                    }
                    umma2_commit_both(&empty[s]);
                }
                if (!ok) break;
                umma2_commit_both(&tmem_full[acc]);
            }
            if (!ok) atomicExch(g.err, 212);
        }
    } else {
        // ================= epilogue warps (both CTAs): this CTA's 128 rows x BNT columns =================
        // Output leaves through a per-warp staging buffer: the lane that owns a row writes its 16-column chunks into the
        // buffer, then the warp stores whole 128-byte row segments (8 lanes per row, 4 rows per instruction).  Straight from
        // registers every store instruction touched 32 different lines, and the tile's stores cost more than its MMAs' loads.
        // Everything that does not need the accumulator is done BEFORE waiting for it, while the tile's MMAs run: the
        // dropout keep bits (EPI_FWD: Philox) and the gate bits (EPI_DX: the gate rows come in through the staging buffer
        // as whole row segments) are packed one bit per column into `bits`.
        constexpr bool OUT_BF16 = (EPI == EPI_FWD || EPI == EPI_DX);
        constexpr int SPAN_COLS = OUT_BF16 ? 64 : 32;            // 128 bytes of an output row
        constexpr int SPAN_CHUNKS = SPAN_COLS / 16;
        constexpr int WARP_COLS = BNT / (EW / 4);               // EW / 4 warps share a 32-lane quarter of the accumulator
        const int q = warp & 3;
        const int half = (warp - 3) >> 2;
        const uint32_t stg = wxtc::smem_u32(sEpi + size_t(warp - 3) * EPI_STAGE_BYTES);
        const int srow = lane >> 3, sunit = lane & 7;            // store phase: 4 rows per instruction, 8 units per row
        uint32_t local = 0;
        for (int64_t tile = cl; tile < n_tiles; tile += n_cl, ++local) {
            const uint32_t acc = local % NACC, use = local / NACC;
            int64_t mn, out_off;
            int kb0, kb1;
            decode(tile, mn, kb0, kb1, out_off);
            const int64_t m0 = (mn / tiles_n) * 2 * BM + int64_t(rank) * BM, n0 = (mn % tiles_n) * BNT;
            const int64_t row0 = m0 + q * 32;
            const int64_t row = row0 + lane;
            const bool row_ok = row < g.M;
            const int wcol0 = half * WARP_COLS;                   // this warp's first column within the tile

            // one bit per column of this lane's row, kept as a queue: words are pushed at the top in column order here and
            // consumed from the bottom span by span below (constant register indices without unrolling the loops)
            constexpr int NW = WARP_COLS / 32;
            uint32_t bits[NW];
#pragma unroll
            for (int i = 0; i < NW; ++i) bits[i] = 0xFFFFFFFFu;
            if (EPI == EPI_FWD && g.keep_thresh16 < 65536u) {
#pragma unroll 1
                for (int i = 0; i < NW; ++i) {
                    uint32_t word = 0;
#pragma unroll 1
                    for (int h = 0; h < 4; ++h) {
                        uint32_t r[4];
                        philox4(uint32_t(row), uint32_t(((n0 + wcol0) >> 3) + i * 4 + h), g.layer, g.step, g.seed_lo, g.seed_hi, r);
                        uint32_t b8 = 0;
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const uint32_t u = (r[e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
                            b8 |= uint32_t(u < g.keep_thresh16) << e;
                        }
                        word |= b8 << (8 * h);
                    }
#pragma unroll
                    for (int k = 0; k + 1 < NW; ++k) bits[k] = bits[k + 1];
                    bits[NW - 1] = word;
                }
            }
            if (EPI == EPI_DX) {
#pragma unroll 1
                for (int sp = 0; sp < WARP_COLS / 64; ++sp) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int r = 4 * i + srow;
                        uint4 gv = make_uint4(0, 0, 0, 0);
                        if (row0 + r < g.M)
                            gv = __ldg(reinterpret_cast<const uint4*>(g.gate + (row0 + r) * g.ldg + n0 + wcol0 + sp * 64) + sunit);
                        sts128(epi_unit_addr(stg, r, sunit), gv.x, gv.y, gv.z, gv.w);
                    }
                    __syncwarp();
                    uint32_t w01[2] = {0, 0};
#pragma unroll
                    for (int hw = 0; hw < 2; ++hw) {
#pragma unroll 1
                        for (int u = 0; u < 4; ++u) {
                            const uint4 gq = lds128(epi_unit_addr(stg, lane, hw * 4 + u));
                            const uint32_t gw[4] = {gq.x, gq.y, gq.z, gq.w};
                            uint32_t b8 = 0;
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                const uint32_t hb = (gw[e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
                                b8 |= uint32_t((hb & 0x7FFFu) != 0 && !(hb & 0x8000u)) << e;      // gate > 0
                            }
                            w01[hw] |= b8 << (8 * u);
                        }
                    }
#pragma unroll
                    for (int k = 0; k + 2 < NW; ++k) bits[k] = bits[k + 2];
                    bits[NW - 2] = w01[0];
                    bits[NW - 1] = w01[1];
                    __syncwarp();
                }
            }

            if (!wxtc::mbar_wait(&tmem_full[acc], use & 1)) {
                if (lane == 0) atomicExch(g.err, 213);
                break;
            }
            wxtc::tc_fence_after();
#pragma unroll 1
            for (int sp = 0; sp < WARP_COLS / SPAN_COLS; ++sp) {
                const int cbase = wcol0 + sp * SPAN_COLS;                     // first column of the span within the tile
                float v[SPAN_CHUNKS][16];
#pragma unroll
                for (int ch = 0; ch < SPAN_CHUNKS; ++ch)
                    wxtc::tmem_ld16(tmem_base + (uint32_t(q * 32) << 16) + acc * BN + cbase + ch * 16, v[ch]);
                wxtc::tmem_ld_wait();
#pragma unroll
                for (int ch = 0; ch < SPAN_CHUNKS; ++ch) {
                    float (&x)[16] = v[ch];
                    const uint32_t b16 = (bits[ch >> 1] >> (16 * (ch & 1))) & 0xFFFFu;     // bottom of the queue = this span
                    if (EPI == EPI_FWD || EPI == EPI_PLAIN) {
                        if (g.bias != nullptr) {
                            const float4* bp = reinterpret_cast<const float4*>(g.bias + n0 + cbase + ch * 16);
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const float4 b = __ldg(bp + k);
                                x[4 * k] += b.x; x[4 * k + 1] += b.y; x[4 * k + 2] += b.z; x[4 * k + 3] += b.w;
                            }
                        }
                    }
                    if (EPI == EPI_FWD) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) x[i] = ((b16 >> i) & 1u) ? fmaxf(x[i], 0.0f) * g.inv_keep : 0.0f;
                    }
                    if (EPI == EPI_DX) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) x[i] = (row_ok && ((b16 >> i) & 1u)) ? x[i] * g.inv_keep : 0.0f;
                    }
                    if (OUT_BF16) {
                        sts128(epi_unit_addr(stg, lane, 2 * ch), pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]),
                               pack_bf16(x[4], x[5]), pack_bf16(x[6], x[7]));
                        sts128(epi_unit_addr(stg, lane, 2 * ch + 1), pack_bf16(x[8], x[9]), pack_bf16(x[10], x[11]),
                               pack_bf16(x[12], x[13]), pack_bf16(x[14], x[15]));
                    } else {
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            sts128(epi_unit_addr(stg, lane, 4 * ch + u), __float_as_uint(x[4 * u]), __float_as_uint(x[4 * u + 1]),
                                   __float_as_uint(x[4 * u + 2]), __float_as_uint(x[4 * u + 3]));
                    }
                }
                __syncwarp();
                {
                    unsigned char* obase = OUT_BF16
                        ? reinterpret_cast<unsigned char*>(reinterpret_cast<__nv_bfloat16*>(g.out) + n0 + cbase)
                        : reinterpret_cast<unsigned char*>(reinterpret_cast<float*>(g.out) + out_off + n0 + cbase);
                    const int64_t pitch = g.ldo * (OUT_BF16 ? 2 : 4);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int r = 4 * i + srow;
                        const uint4 o = lds128(epi_unit_addr(stg, r, sunit));
                        if (row0 + r < g.M)
                            *reinterpret_cast<uint4*>(obase + (row0 + r) * pitch + sunit * 16) = o;
                    }
                }
                if (EPI == EPI_DX && g.colsum_part != nullptr) {
                    // column sums over this warp's 32 rows of the values AS STORED (bf16): lane l owns columns 2l, 2l+1 of the span
                    float s0[4] = {0.0f, 0.0f, 0.0f, 0.0f}, s1[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                    for (int r = 0; r < 32; ++r) {
                        uint32_t wv;
                        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(wv) : "r"(epi_unit_addr(stg, r, lane >> 2) + (lane & 3) * 4) : "memory");
                        s0[r & 3] += __uint_as_float(wv << 16);
                        s1[r & 3] += __uint_as_float(wv & 0xFFFF0000u);
                    }
                    float2 cs;
                    cs.x = (s0[0] + s0[1]) + (s0[2] + s0[3]);
                    cs.y = (s1[0] + s1[1]) + (s1[2] + s1[3]);
                    *reinterpret_cast<float2*>(g.colsum_part + (row0 >> 5) * g.N + n0 + cbase + 2 * lane) = cs;
                }
                if (OUT_BF16) {                                               // pop the two words this span used
#pragma unroll
                    for (int k = 0; k + 2 < NW; ++k) bits[k] = bits[k + 2];
                }
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
