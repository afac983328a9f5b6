// Tensor-core (tcgen05) version of the fused weather-MLP step: [standardise +] forward + cross-entropy + backward
// (+ Adam).  Replaces /root/reference/jobs/train_lightning_ddp.py:57-71,:88 (and, with the raw-column loader,
// /root/reference/jobs/preprocess.py:38-41) like mlp_grid_kernel, but the two F x 64 contractions run on the
// 5th-generation tensor cores:
//
//   GEMM1   z[128 x 64]     = Xaug[128 x K1] . W1aug[64 x K1]^T      (M=128, N=64, K1 = ceil((F+1)/8)*8, K-major, SW128)
//   GEMM2   G[64 x N2]     += DH[128 x 64]^T . Xaug[128 x N2]        (M=64,  N2 = ceil((F+3)/8)*8, K=128 rows, MN-major, SW128_B32)
//
// with Xaug = [x_0..x_{F-1}, 1, dlogit_0, dlogit_1, 0...]: the ones column turns b1 into a GEMM1 weight and db1 into a
// GEMM2 output; G accumulates in TMEM over ALL tiles of the CTA (dW1 | db1 never leave tensor memory until the end).
// fp32 accuracy is kept with a 3-term tf32 split of both operands (hi*lo + lo*hi + hi*hi, ~2^-21 relative), so the
// result meets the same 1e-5 parity bound as the FFMA kernel.
//
// Roles (320 threads, one CTA per SM, persistent over 128-row tiles; tile i belongs to worker group i & 1):
//   warps 0-3 / 4-7  two row-worker groups: thread r <-> tile row r <-> TMEM lane r.  A group turns the TMA-landed rows
//              of its tile into the swizzled hi/lo operand tiles (standardising raw columns on the way), reads z from
//              TMEM, does ReLU / dropout / logits / softmax / dlogits / dh in registers (row-local: no shuffles),
//              writes DH and Xaug for GEMM2, and keeps its dW2 sums in its own TMEM columns (read-modify-write),
//              db2 / loss sums in registers.  While one group is in its pointwise phase the other converts / waits.
//   warp 8     TMA producer: cp.async.bulk of the tile's rows (or F column segments), labels and dropout keep bits
//              into a 3-stage ring.
//   warp 9     MMA issuer: one elected thread issues tcgen05.mma and commits to mbarriers.
#include <math.h>
#include <string.h>

#include "mlp_kernels.cuh"
#include "tc_common.cuh"

namespace {
using namespace wxtc;
using wxk::AdamScalars;
using wxk::GridArgs;

constexpr int TILE = 128;
constexpr int NGROUPS = 2;
constexpr int WORKERS = NGROUPS * TILE;          // 256 row-worker threads
constexpr int TC_THREADS = WORKERS + 64;         // + producer warp + MMA warp
constexpr int PRODUCER_WARP = WORKERS / 32, MMA_WARP = WORKERS / 32 + 1;
constexpr int STAGES = 3;
constexpr int MAXF_TC = 29;                      // F feature columns + 3 fixed columns fill the 32-column operand tile
constexpr int COL_ONE = 29, COL_DL0 = 30, COL_DL1 = 31;   // fixed positions: selects on them are compile-time
constexpr int TMEM_COLS = 512;                   // D1 x2 [0,128) | G [128,160) | dW2 sums of group g [256+128g, 384+128g)
constexpr int TM_G = 128, TM_ACC = 256;

struct TcSmem {
    // operand tiles (all 1024-byte aligned)
    float w_hi[64 * 32], w_lo[64 * 32];          // W1aug [n][32]   K-major SW128
    float xk_hi[TILE * 32], xk_lo[TILE * 32];    // Xaug  [row][32] K-major SW128      (GEMM1 A)
    float xm_hi[TILE * 32], xm_lo[TILE * 32];    // Xaug  [row][32] MN-major SW128_B32 (GEMM2 B)
    float a_hi[64 * TILE], a_lo[64 * TILE];      // DH    [k=row][m=64]  MN-major SW128_B32 (GEMM2 A); reused for the final dW2 reduce
    // TMA landing ring: dense rows [128][F] (row-major loader) or column segments [F][128] (raw loader)
    float land_x[STAGES][TILE * MAXF_TC];
    long long land_y[STAGES][TILE];
    unsigned long long land_m[STAGES][TILE];
    __align__(16) float w2[2 * 64 + 4];          // W2[c][j], b2[0..1]
    double mean[32], rstd[32];                   // raw loader: column mean and 1/std_val
    float slab[64 * 32 + 64 + 128 + 2 + 4];      // CTA partial in flat-parameter order (+ loss)
    float red_small[WORKERS][4];                 // db2_0, db2_1, loss per row worker
    uint64_t land_full[STAGES], land_empty[STAGES];
    uint64_t xk_ready, a2_ready, a2_free, d1_full[2];
    uint32_t tmem_base;
    int err;
    bool last;
    AdamScalars adam;
};

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(void* base, uint32_t off, float a, float b, float c, float d) {
    *reinterpret_cast<float4*>(static_cast<unsigned char*>(base) + off) = make_float4(a, b, c, d);
}

// RAW: raw-column loader (column segments [F][128] in the landing buffer, standardised on the way) vs dense row-major rows.
// A template parameter so that every landing-buffer access has a compile-time stride (immediate LDS offsets).
template <bool RAW>
__global__ void __launch_bounds__(TC_THREADS, 1) tc_step_kernel(const GridArgs a, const int F, int* __restrict__ err_out) {
    extern __shared__ __align__(1024) unsigned char raw[];     // 1024-byte alignment: the swizzle atoms need it
    TcSmem& s = *reinterpret_cast<TcSmem*>(raw);               // direct cast keeps the shared address space (LDS/STS)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int P = 64 * F + 194;
    const int OFF_B1 = 64 * F, OFF_W2 = 64 * F + 64, OFF_B2 = 64 * F + 192;
    const int slab_stride = (P + 1 + 3) & ~3;
    constexpr int K1 = 32;                        // GEMM1 K extent: features [0,F) | zeros | 1 at column 29 (| dl at 30,31: zero weights)
    constexpr int N2 = 32;                        // GEMM2 N extent
    constexpr bool raw_cols = RAW;
    const int64_t ntiles = (a.batch + TILE - 1) / TILE;
    const int64_t my_tiles = (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x;     // >= 1 by construction of the grid

    if (tid == 0 && (smem_u32(raw) & 1023u) != 0) atomicExch(err_out, 99);   // dynamic smem base not 1024-aligned
    if (a.dbg != nullptr && tid == 0) {
        unsigned smid;
        long long t;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.dbg[blockIdx.x * 4 + 0] = smid;
        a.dbg[blockIdx.x * 4 + 1] = t;
    }
    if (tid == 0) {
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(&s.land_full[i], 1);
            mbar_init(&s.land_empty[i], TILE);
        }
        mbar_init(&s.xk_ready, TILE);
        mbar_init(&s.a2_ready, TILE);
        mbar_init(&s.a2_free, 1);
        mbar_init(&s.d1_full[0], 1);
        mbar_init(&s.d1_full[1], 1);
        fence_mbar_init();
        s.err = 0;
    }
    if (warp == 0) tmem_alloc(&s.tmem_base, TMEM_COLS);
    // W1aug operand (hi/lo, K-major SW128): row n = [W1[n][0..F-1], b1[n], 0 ...]; W2 / b2 / statistics to smem
    if (tid < 64) {
        const int n = tid;
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
            float hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int c = c4 * 4 + e;
                const float v = c < F ? a.params[n * F + c] : (c == COL_ONE ? a.params[OFF_B1 + n] : 0.0f);
                split_tf32(v, hi[e], lo[e]);
            }
            const uint32_t off = sw128_offset(n, c4 * 4);
            st4(s.w_hi, off, hi[0], hi[1], hi[2], hi[3]);
            st4(s.w_lo, off, lo[0], lo[1], lo[2], lo[3]);
        }
    }
    for (int i = tid; i < 130; i += TC_THREADS) s.w2[i] = a.params[OFF_W2 + i];     // W2[2][64] then b2[2] (contiguous)
    if (raw_cols && tid < F) {
        s.mean[tid] = a.stats[tid * 4 + 1];
        s.rstd[tid] = 1.0 / a.stats[tid * 4 + 3];
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = s.tmem_base;
    const uint32_t d1_base = tm;
    const uint32_t g_base = tm + TM_G;
    const uint64_t step_word = (a.state != nullptr) ? uint64_t(a.state[0]) : a.step;

    auto tile_rows = [&](int64_t t) -> int {
        const int64_t left = a.batch - t * TILE;
        return left < TILE ? int(left) : TILE;
    };

    if (warp == PRODUCER_WARP) {
        // ================= TMA producer =================
        for (int64_t i = 0; i < my_tiles; ++i) {
            const int st = int(i % STAGES);
            const int64_t tile = blockIdx.x + i * gridDim.x;
            if (i >= STAGES && !mbar_wait(&s.land_empty[st], uint32_t((i / STAGES - 1) & 1))) s.err = 2;
            const int nr = tile_rows(tile);
            const int64_t row = a.row0 + tile * TILE;
            const bool has_m = a.dropout_mode == WX_DROPOUT_MASK;
            if (nr == TILE) {           // full tile: every size is a multiple of 16 B; start alignment is checked on the host
                const uint32_t by = TILE * 8;
                if (lane == 0) mbar_arrive_expect_tx(&s.land_full[st], uint32_t(TILE) * F * 4 + by + (has_m ? by : 0u));
                __syncwarp();
                if (!raw_cols) {
                    if (lane == 0) bulk_g2s(s.land_x[st], a.x + row * F, uint32_t(TILE) * F * 4, &s.land_full[st]);
                } else if (lane < F) {  // one 512-byte column segment per lane
                    bulk_g2s(s.land_x[st] + lane * TILE, static_cast<const float*>(a.cols.col[lane]) + row, TILE * 4,
                             &s.land_full[st]);
                }
                if (lane == 0) {
                    bulk_g2s(s.land_y[st], a.y + row, by, &s.land_full[st]);
                    if (has_m) bulk_g2s(s.land_m[st], a.keep_mask + tile * TILE, by, &s.land_full[st]);
                }
            } else {
                // ragged last tile: plain loads by the whole warp, then a normal arrive
                if (!raw_cols) {
                    const float* gx = a.x + row * F;
                    for (int e = lane; e < nr * F; e += 32) s.land_x[st][e] = gx[e];
                } else {
                    for (int c = 0; c < F; ++c)
                        for (int e = lane; e < nr; e += 32)
                            s.land_x[st][c * TILE + e] = static_cast<const float*>(a.cols.col[c])[row + e];
                }
                for (int e = lane; e < nr; e += 32) {
                    s.land_y[st][e] = a.y[row + e];
                    if (has_m) s.land_m[st][e] = a.keep_mask[tile * TILE + e];
                }
                __threadfence_block();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s.land_full[st]);
            }
        }
    } else if (warp == MMA_WARP) {
        // ================= MMA issuer =================
        {
            // The WHOLE warp runs this code and one elected lane issues each MMA: operands stay warp-uniform (uniform
            // registers, no per-instruction conversion) -- 3x faster issue than a lone thread (profiles/: 84 -> 24 cycles
            // per MMA).  One base descriptor per operand/term; a k-step only adds a constant to the start-address field.
            const uint32_t id1 = make_idesc_tf32(128, 64, false, false);
            const uint32_t id2 = make_idesc_tf32(64, N2, true, true);
            const uint64_t xk_d[2] = {make_smem_desc(smem_u32(s.xk_hi), 16, 1024), make_smem_desc(smem_u32(s.xk_lo), 16, 1024)};
            const uint64_t wk_d[2] = {make_smem_desc(smem_u32(s.w_hi), 16, 1024), make_smem_desc(smem_u32(s.w_lo), 16, 1024)};
            const uint64_t am_d[2] = {make_smem_desc(smem_u32(s.a_hi), 512, 1024, LAYOUT_SW128_B32),
                                      make_smem_desc(smem_u32(s.a_lo), 512, 1024, LAYOUT_SW128_B32)};
            const uint64_t xm_d[2] = {make_smem_desc(smem_u32(s.xm_hi), 512, 512, LAYOUT_SW128_B32),
                                      make_smem_desc(smem_u32(s.xm_lo), 512, 512, LAYOUT_SW128_B32)};
            const int k1steps = K1 / 8;
            auto gemm1 = [&](uint32_t d1) {
                // small terms first: hi*lo, lo*hi, then hi*hi
                bool acc = false;
#pragma unroll
                for (int term = 0; term < 3; ++term) {
                    const uint64_t ad = xk_d[term == 1 ? 1 : 0], bd = wk_d[term == 0 ? 1 : 0];
#pragma unroll 4
                    for (int k = 0; k < k1steps; ++k) {
                        umma_tf32_warp(d1, ad + uint64_t(k * 2), bd + uint64_t(k * 2), id1, acc);       // +32 B per k-step
                        acc = true;
                    }
                }
            };
            auto gemm2 = [&](bool first) {
                bool acc = !first;
#pragma unroll
                for (int term = 0; term < 3; ++term) {
                    const uint64_t ad = am_d[term == 1 ? 1 : 0], bd = xm_d[term == 0 ? 1 : 0];
#pragma unroll
                    for (int k = 0; k < TILE / 8; ++k) {
                        umma_tf32_warp(g_base, ad + uint64_t(k * 128), bd + uint64_t(k * 64), id2, acc);  // +2048 B / +1024 B
                        acc = true;
                    }
                }
            };
            const bool mt = a.dbg != nullptr && lane == 0;
            long long mph[6] = {0, 0, 0, 0, 0, 0};
            long long mk = mt ? clock64() : 0;
#define WX_MPH(k)                               \
    do {                                        \
        if (mt) {                               \
            const long long now__ = clock64();  \
            mph[k] += now__ - mk;               \
            mk = now__;                         \
        }                                       \
    } while (0)
            for (int64_t i = 0; i < my_tiles; ++i) {
                if (!mbar_wait_hot(&s.xk_ready, uint32_t(i & 1))) s.err = 3;
                tc_fence_after();
                WX_MPH(0);
                gemm1(d1_base + uint32_t(i & 1) * 64);
                umma_commit_warp(&s.d1_full[i & 1]);
                WX_MPH(1);
                if (mt && (i & 7) == 3) {          // sampled: how long does GEMM1 take to complete once issued?
                    mbar_wait_hot(&s.d1_full[i & 1], uint32_t((i >> 1) & 1));
                    WX_MPH(4);
                }
                if (i > 0) {
                    if (!mbar_wait_hot(&s.a2_ready, uint32_t((i - 1) & 1))) s.err = 4;
                    tc_fence_after();
                    WX_MPH(2);
                    gemm2(i - 1 == 0);
                    umma_commit_warp(&s.a2_free);
                    WX_MPH(3);
                    if (mt && (i & 7) == 5) {      // sampled: GEMM2 completion latency
                        mbar_wait_hot(&s.a2_free, uint32_t((i - 1) & 1));
                        WX_MPH(5);
                    }
                }
            }
            if (mt)
                for (int k = 0; k < 6; ++k) a.dbg[1184 * 4 + 148 * 16 + blockIdx.x * 8 + k] = mph[k];
            if (!mbar_wait_hot(&s.a2_ready, uint32_t((my_tiles - 1) & 1))) s.err = 4;
            tc_fence_after();
            gemm2(my_tiles - 1 == 0);
            umma_commit_warp(&s.a2_free);
        }
    } else {
        // ================= row workers: group 0 = warps 0-3, group 1 = warps 4-7 =================
        const int grp = warp >> 2;
        const int r = tid & (TILE - 1);                           // tile row == TMEM lane
        const uint32_t lane_addr = uint32_t((warp & 3) * 32) << 16;
        const uint32_t acc_base = tm + TM_ACC + uint32_t(grp) * 128 + lane_addr;
        const bool drop = a.dropout_mode != WX_DROPOUT_OFF;
        const float ds = drop ? 1.25f : 1.0f;                     // 1/(1-0.2), train_lightning_ddp.py:60
        const uint2 key = make_uint2(uint32_t(a.seed), uint32_t(a.seed >> 32));
        float gb2_0 = 0.0f, gb2_1 = 0.0f, loss_acc = 0.0f;
        const float b2_0 = s.w2[128], b2_1 = s.w2[129];
        // optional phase timers (wx_debug_cta_times): cycles spent per phase by row 0 of each group, summed over tiles
        const bool timing = a.dbg != nullptr && r == 0;
        long long tph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        long long tmark = timing ? clock64() : 0;
#define WX_TPH(k)                               \
    do {                                        \
        if (timing) {                           \
            const long long now__ = clock64();  \
            tph[k] += now__ - tmark;            \
            tmark = now__;                      \
        }                                       \
    } while (0)
        {   // zero this thread's dW2 sums in TMEM
            float zero[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) zero[q] = 0.0f;
#pragma unroll
            for (int c = 0; c < 8; ++c) tmem_st16(acc_base + c * 16, zero);
            tmem_st_wait();
        }

        // Xaug row of tile-slot `i` -> K-major hi/lo operand tile (GEMM1 A).  With the raw loader the row is
        // standardised here, ((double)x - mean) * (1/std) rounded to fp32 (preprocess.py:38-41 + the cast at
        // train_lightning_ddp.py:45, to within 1 ulp of the exact division), and written back to the landing buffer
        // for its second use in phase B.
        auto convert = [&](int64_t i) {
            const int st = int(i % STAGES);
            const int64_t tile = blockIdx.x + i * gridDim.x;
            if (!mbar_wait(&s.land_full[st], uint32_t((i / STAGES) & 1))) s.err = 5;
            const bool valid = r < tile_rows(tile);
            float* lx = s.land_x[st] + (raw_cols ? r : r * F);     // element c of this row at lx[c * cstride]
            constexpr int cstride = raw_cols ? TILE : 1;
#pragma unroll 1
            for (int c8 = 0; c8 < 4; ++c8) {
                float t[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    // unconditional loads at compile-time offsets, all eight in flight; columns >= F read whatever follows in
                    // the landing ring (still inside TcSmem) and are discarded by the selects below
                    const int c = c8 * 8 + e;
                    t[e] = lx[c * cstride];
                }
                if (raw_cols) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const int c = c8 * 8 + e;
                        if (c < F) {
                            const float xr = t[e];
                            float v = __double2float_rn((double(xr) - s.mean[c]) * s.rstd[c]);
                            if (xr != xr) v = (a.impute == WX_IMPUTE_MEAN) ? 0.0f : __int_as_float(0x7fc00000);
                            t[e] = v;
                            lx[c * cstride] = v;                    // second use in phase B
                        }
                    }
                }
                float hi[8], lo[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const int c = c8 * 8 + e;
                    const float v = !valid ? 0.0f : (c == COL_ONE ? 1.0f : (c < COL_ONE && c < F ? t[e] : 0.0f));
                    split_tf32(v, hi[e], lo[e]);
                }
                const uint32_t off0 = sw128_offset(r, c8 * 8), off1 = sw128_offset(r, c8 * 8 + 4);
                st4(s.xk_hi, off0, hi[0], hi[1], hi[2], hi[3]);
                st4(s.xk_hi, off1, hi[4], hi[5], hi[6], hi[7]);
                st4(s.xk_lo, off0, lo[0], lo[1], lo[2], lo[3]);
                st4(s.xk_lo, off1, lo[4], lo[5], lo[6], lo[7]);
            }
            fence_proxy_async();
            mbar_arrive(&s.xk_ready);
        };

        // One convert call site (keeps the code small): each pass first converts this group's next tile-slot -- slot 0
        // at once, any later slot once XK is back, i.e. GEMM1 of the slot before it has completed -- then processes it.
        for (int64_t i = grp;; i += NGROUPS) {
            if (i >= my_tiles) break;
            if (i > 0 && !mbar_wait(&s.d1_full[(i - 1) & 1], uint32_t(((i - 1) >> 1) & 1))) s.err = 10;
            convert(i);
            WX_TPH(6);
            const int st = int(i % STAGES);
            const int64_t tile = blockIdx.x + i * gridDim.x;
            const bool valid = r < tile_rows(tile);
            const uint32_t d1 = d1_base + uint32_t(i & 1) * 64 + lane_addr;
            WX_TPH(7);
            // ---- dropout keep bits for this row: independent of GEMM1, so they are generated while it runs ----
            uint64_t keep = ~0ull;
            if (a.dropout_mode == WX_DROPOUT_MASK) keep = s.land_m[st][r];
            else if (a.dropout_mode == WX_DROPOUT_PHILOX) keep = wxk::philox_keep_mask(uint64_t(tile * TILE + r), key, step_word);
            const int yv = int(s.land_y[st][r]);
            WX_TPH(1);
            if (!mbar_wait(&s.d1_full[i & 1], uint32_t((i >> 1) & 1))) s.err = 6;
            tc_fence_after();
            WX_TPH(0);

            // ---- phase A: a = dropout(relu(z)), logits ----
            float lg0 = 0.0f, lg1 = 0.0f, lg0b = 0.0f, lg1b = 0.0f;
            uint64_t gate = 0;
#pragma unroll 1
            for (int cc = 0; cc < 4; ++cc) {
                float z[16];
                tmem_ld16(d1 + cc * 16, z);
                float w0[16], w1[16];
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                    const float4 u = ld4(&s.w2[cc * 16 + q4 * 4]), t = ld4(&s.w2[64 + cc * 16 + q4 * 4]);
                    w0[q4 * 4] = u.x; w0[q4 * 4 + 1] = u.y; w0[q4 * 4 + 2] = u.z; w0[q4 * 4 + 3] = u.w;
                    w1[q4 * 4] = t.x; w1[q4 * 4 + 1] = t.y; w1[q4 * 4 + 2] = t.z; w1[q4 * 4 + 3] = t.w;
                }
                const uint32_t kb = uint32_t(keep >> (cc * 16)) & 0xffffu;
                uint32_t gb = 0;
                tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const float h = (z[q] <= 0.0f) ? 0.0f : z[q];                     // relu, NaN passes like torch
                    const float av = h * (((kb >> q) & 1u) ? ds : 0.0f);
                    if (q & 1) {
                        lg0b = fmaf(av, w0[q], lg0b);
                        lg1b = fmaf(av, w1[q], lg1b);
                    } else {
                        lg0 = fmaf(av, w0[q], lg0);
                        lg1 = fmaf(av, w1[q], lg1);
                    }
                    gb |= uint32_t(!(av <= 0.0f)) << q;
                }
                gate |= uint64_t(gb) << (cc * 16);
            }
            lg0 = (lg0 + lg0b) + b2_0;
            lg1 = (lg1 + lg1b) + b2_1;
            // ---- softmax / loss / dlogits, ATen's operation order ----
            const float mx = fmaxf(lg0, lg1);
            const float s0 = lg0 - mx, s1 = lg1 - mx;
            const float lsum = logf(expf(s0) + expf(s1));
            const float lp0 = s0 - lsum, lp1 = s1 - lsum;
            float dl0 = 0.0f, dl1 = 0.0f;
            if (valid) {
                loss_acc += -(yv ? lp1 : lp0);
                dl0 = (expf(lp0) - (yv == 0 ? 1.0f : 0.0f)) * a.scale;
                dl1 = (expf(lp1) - (yv == 1 ? 1.0f : 0.0f)) * a.scale;
                gb2_0 += dl0;
                gb2_1 += dl1;
            }
            WX_TPH(2);
            // ---- phase B (still off the critical path: touches registers and TMEM only) ----
            // dW2 sums (TMEM read-modify-write) and dh for all 64 units, kept in registers until the GEMM2 operand
            // buffers are free.  ds is folded into the dlogits: (dl*ds)*zg == dl*(zg*ds) up to one rounding.
            float dhv[64];
            {
                const float dl0s = dl0 * ds, dl1s = dl1 * ds;
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) {
                    float z[16], g0[16], g1[16];
                    tmem_ld16(d1 + cc * 16, z);
                    tmem_ld16(acc_base + cc * 16, g0);
                    tmem_ld16(acc_base + 64 + cc * 16, g1);
                    const uint32_t gb = uint32_t(gate >> (cc * 16)) & 0xffffu;
                    float w0[16], w1[16];
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        const float4 u = ld4(&s.w2[cc * 16 + q4 * 4]), t = ld4(&s.w2[64 + cc * 16 + q4 * 4]);
                        w0[q4 * 4] = u.x; w0[q4 * 4 + 1] = u.y; w0[q4 * 4 + 2] = u.z; w0[q4 * 4 + 3] = u.w;
                        w1[q4 * 4] = t.x; w1[q4 * 4 + 1] = t.y; w1[q4 * 4 + 2] = t.z; w1[q4 * 4 + 3] = t.w;
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const bool on = (gb >> q) & 1u;                    // unit active: z > 0 (or NaN) and kept
                        const float zg = on ? z[q] : 0.0f;
                        g0[q] = fmaf(dl0s, zg, g0[q]);
                        g1[q] = fmaf(dl1s, zg, g1[q]);
                        dhv[cc * 16 + q] = on ? fmaf(dl1s, w1[q], dl0s * w0[q]) : 0.0f;
                    }
                    tmem_st16(acc_base + cc * 16, g0);
                    tmem_st16(acc_base + 64 + cc * 16, g1);
                }
            }
            // ---- GEMM2 operands may be rewritten once the previous tile's GEMM2 has drained ----
            if (i > 0 && !mbar_wait(&s.a2_free, uint32_t((i - 1) & 1))) s.err = 7;
            WX_TPH(3);
            // Xaug row (MN-major): features, 1, dl0, dl1
            {
                const float* lx = s.land_x[st] + (raw_cols ? r : r * F);
                constexpr int cstride = raw_cols ? TILE : 1;
#pragma unroll 1
                for (int g8 = 0; g8 < 4; ++g8) {
                    float t[8], hi[8], lo[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const int c = g8 * 8 + e;
                        t[e] = lx[c * cstride];
                    }
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const int c = g8 * 8 + e;
                        const float v = !valid ? 0.0f
                                               : (c == COL_ONE ? 1.0f : (c == COL_DL0 ? dl0 : (c == COL_DL1 ? dl1 : (c < F ? t[e] : 0.0f))));
                        split_tf32(v, hi[e], lo[e]);
                    }
                    const uint32_t off = mn32_offset(g8 * 8, r, 512, 512);
                    st4(s.xm_hi, off, hi[0], hi[1], hi[2], hi[3]);
                    st4(s.xm_hi, off + 16, hi[4], hi[5], hi[6], hi[7]);
                    st4(s.xm_lo, off, lo[0], lo[1], lo[2], lo[3]);
                    st4(s.xm_lo, off + 16, lo[4], lo[5], lo[6], lo[7]);
                }
            }
            WX_TPH(4);
            // ---- critical section (serialised between the two groups by a2_free): operand stores only ----
#pragma unroll
            for (int g8 = 0; g8 < 8; ++g8) {
                float hi[8], lo[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) split_tf32(dhv[g8 * 8 + e], hi[e], lo[e]);
                const uint32_t off = mn32_offset(g8 * 8, r, 512, 1024);
                st4(s.a_hi, off, hi[0], hi[1], hi[2], hi[3]);
                st4(s.a_hi, off + 16, hi[4], hi[5], hi[6], hi[7]);
                st4(s.a_lo, off, lo[0], lo[1], lo[2], lo[3]);
                st4(s.a_lo, off + 16, lo[4], lo[5], lo[6], lo[7]);
            }
            tmem_st_wait();
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(&s.a2_ready);
            mbar_arrive(&s.land_empty[st]);
            WX_TPH(5);
        }
        if (timing)
            for (int k = 0; k < 8; ++k) a.dbg[1184 * 4 + (blockIdx.x * 2 + grp) * 8 + k] = tph[k];
        // ---- all tiles done: wait for the last GEMM2, then assemble the CTA slab ----
        if (!mbar_wait(&s.a2_free, uint32_t((my_tiles - 1) & 1))) s.err = 8;
        tc_fence_after();
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (grp == 0) {
            // G[m][n] (M=64 accumulator: row m = 16*warp + lane, lanes 0-15 of each sub-partition)
            for (int c0 = 0; c0 < N2; c0 += 16) {
                float v[16];
                tmem_ld16(g_base + lane_addr + c0, v);
                tmem_ld_wait();
                if (lane < 16) {
                    const int m = (warp & 3) * 16 + lane;
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const int n = c0 + q;
                        if (n < F) s.slab[m * F + n] = v[q];             // dW1[m][n]
                        else if (n == COL_ONE) s.slab[OFF_B1 + m] = v[q];  // db1[m]
                    }
                }
            }
        }
        // dW2 / db2 / loss: per-thread sums -> ordered column sums through smem (a_hi/a_lo are free now), one group at a time
        float* red = s.a_hi;                                          // [128 rows][128] floats = 64 KB = a_hi + a_lo
        s.red_small[tid][0] = gb2_0;
        s.red_small[tid][1] = gb2_1;
        s.red_small[tid][2] = loss_acc;
        float colsum = 0.0f;
#pragma unroll 1
        for (int g = 0; g < NGROUPS; ++g) {
            if (grp == g) {
#pragma unroll 1
                for (int c = 0; c < 8; ++c) {
                    float v[16];
                    tmem_ld16(acc_base + c * 16, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4)
                        *reinterpret_cast<float4*>(&red[r * 128 + c * 16 + q4 * 4]) =
                            make_float4(v[q4 * 4], v[q4 * 4 + 1], v[q4 * 4 + 2], v[q4 * 4 + 3]);
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (grp == 0)
                for (int k = 0; k < TILE; ++k) colsum += red[k * 128 + r];
            asm volatile("bar.sync 1, 256;" ::: "memory");
        }
        if (grp == 0) {
            s.slab[OFF_W2 + r] = colsum;
            if (r < 3) {
                float t = 0.0f;
                for (int k = 0; k < WORKERS; ++k) t += s.red_small[k][r];
                s.slab[r < 2 ? OFF_B2 + r : P] = t;
            }
        }
    }
    if (a.dbg != nullptr && tid == 0) {
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.dbg[blockIdx.x * 4 + 2] = t;
        a.dbg[blockIdx.x * 4 + 3] = t;
    }
    // ================= CTA slab -> global, last CTA reduces (same deterministic protocol as mlp_grid_kernel) ==========
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, TMEM_COLS);
    float* mypart = a.part + size_t(blockIdx.x) * slab_stride;
    for (int i = tid; i < P + 1; i += TC_THREADS) mypart[i] = s.slab[i];
    if (tid == 0 && s.err != 0) atomicExch(err_out, s.err);
    __threadfence();
    __syncthreads();
    if (tid == 0) s.last = atomicAdd(a.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s.last) return;
    __threadfence();
    if (a.tail == wxk::TAIL_ADAM && tid == 0) s.adam = wxk::adam_scalars(a.desc, a.state[0] + 1);
    __syncthreads();
    // Deterministic two-level sum of the CTA slabs: work item (float4 group g, segment q of the CTA range) -> partial in
    // shared memory (the operand tiles are dead now), then the SEG partials of a group are added in segment order.
    const unsigned nparts = gridDim.x;
    constexpr int SEG = 4;
    const int ng = slab_stride / 4;
    float4* psum = reinterpret_cast<float4*>(s.xk_hi);            // [SEG][ng] float4 <= 4 * 561 * 16 B = 36 KB (xk_hi + xk_lo)
    const unsigned per = (nparts + SEG - 1) / SEG;
    for (int item = tid; item < ng * SEG; item += TC_THREADS) {
        const int g = item % ng, q = item / ng;
        const unsigned b0 = q * per, b1 = (b0 + per < nparts) ? b0 + per : nparts;
        const float4* src = reinterpret_cast<const float4*>(a.part) + g;
        float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        unsigned b = b0;
        for (; b + 8 <= b1; b += 8) {
            float4 t[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) t[u] = __ldcg(src + size_t(b + u) * ng);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w;
            }
        }
        for (; b < b1; ++b) {
            const float4 t = __ldcg(src + size_t(b) * ng);
            acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
        }
        psum[q * ng + g] = acc;
    }
    __syncthreads();
    for (int i = tid; i < P + 1; i += TC_THREADS) {
        const float* ps = reinterpret_cast<const float*>(psum);
        float sum = 0.0f;
#pragma unroll
        for (int q = 0; q < SEG; ++q) sum += ps[(q * ng) * 4 + i];
        if (i == P) sum *= a.scale;
        s.slab[i] = sum;                                    // this rank's [P+1] bucket
    }
    __syncthreads();
    if (a.peer.world > 1 && a.tail == wxk::TAIL_ADAM) {
        // DDP exchange inside the kernel: one-shot all-reduce over NVLink peer memory (csrc/peer.cuh)
        if (!wxpeer::allreduce_cta(s.slab, P + 1, a.peer, uint64_t(a.state[0]) + 1, tid, TC_THREADS) && tid == 0)
            atomicExch(err_out, 20);
    }
    for (int i = tid; i < P + 1; i += TC_THREADS) {
        const float sum = s.slab[i];
        if (a.tail == wxk::TAIL_GRADS) {
            a.out[i] = sum;
        } else if (i < P) {
            float p = a.p_rw[i], m = a.m[i], v = a.v[i];
            wxk::adam_apply(p, sum, m, v, s.adam);
            a.p_rw[i] = p;
            a.m[i] = m;
            a.v[i] = v;
        } else {
            *a.loss_out = sum;
        }
    }
    if (tid == 0) {
        if (a.dbg != nullptr) {
            long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            a.dbg[blockIdx.x * 4 + 3] = t;
        }
        *a.ticket = 0u;
        if (a.tail == wxk::TAIL_ADAM) a.state[0] += 1;
    }
}

int tc_launch(wx_ctx* ctx, GridArgs& a, cudaStream_t st, const char* who) {
    const int f = ctx->desc.n_features;
    WX_CHECK_ARG(f <= MAXF_TC, "%s: the tensor-core path is built for n_features <= %d (got %d)", who, MAXF_TC, f);
    WX_CHECK_ARG(a.row_idx == nullptr, "%s: the tensor-core path takes contiguous rows only (row_idx must be NULL)", who);
    if (a.loader == wxk::LOAD_ROWMAJOR) {
        WX_CHECK_ARG(((reinterpret_cast<uintptr_t>(a.x) + size_t(a.row0) * f * 4) & 15) == 0,
                     "%s: first row of the batch must start on a 16-byte boundary for the TMA loads", who);
    } else {
        for (int j = 0; j < f; ++j)
            WX_CHECK_ARG(((reinterpret_cast<uintptr_t>(a.cols.col[j]) + size_t(a.row0) * 4) & 15) == 0,
                         "%s: column %d must be 16-byte aligned at row0 for the TMA loads (row0 %% 4 == 0)", who, j);
    }
    WX_CHECK_ARG(((reinterpret_cast<uintptr_t>(a.y) + size_t(a.row0) * 8) & 15) == 0,
                 "%s: labels of the batch must start on a 16-byte boundary for the TMA loads", who);
    WX_CHECK_ARG(a.dropout_mode != WX_DROPOUT_MASK || (reinterpret_cast<uintptr_t>(a.keep_mask) & 15) == 0,
                 "%s: keep_mask must be 16-byte aligned", who);
    WX_CUDA(cudaSetDevice(ctx->device));
    const int smem = int(sizeof(TcSmem));
    // the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute
    static bool attr_set[64] = {};
    if (ctx->device < 0 || ctx->device >= 64 || !attr_set[ctx->device]) {
        WX_CUDA(cudaFuncSetAttribute(tc_step_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        WX_CUDA(cudaFuncSetAttribute(tc_step_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        if (ctx->device >= 0 && ctx->device < 64) attr_set[ctx->device] = true;
    }
    const int64_t ntiles = (a.batch + TILE - 1) / TILE;
    int64_t grid = ntiles < ctx->num_sms ? ntiles : ctx->num_sms;
    const size_t slab = (size_t(64) * f + 194 + 1 + 3) & ~size_t(3);
    if (WX_SCRATCH_PART_OFF + size_t(grid) * slab * 4 > ctx->scratch_bytes) {
        wx_set_error("%s: scratch too small", who);
        return WX_ENOMEM;
    }
    a.part = reinterpret_cast<float*>(ctx->scratch + WX_SCRATCH_PART_OFF);
    a.ticket = reinterpret_cast<unsigned int*>(ctx->scratch) + 1;
    a.desc = ctx->desc;
    a.dbg = ctx->debug_times;
    if (a.tail == wxk::TAIL_ADAM && ctx->peer.world > 1) {
        a.peer = ctx->peer;
        a.scale = a.scale / float(ctx->peer.world);          // mean over W*B rows (what DDP's averaged gradients give)
    }
    int* err = reinterpret_cast<int*>(ctx->scratch) + 8;       // word 8 of the ticket block: sticky protocol-error code
    if (a.loader == wxk::LOAD_ROWMAJOR) tc_step_kernel<false><<<(unsigned)grid, TC_THREADS, smem, st>>>(a, f, err);
    else tc_step_kernel<true><<<(unsigned)grid, TC_THREADS, smem, st>>>(a, f, err);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
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
