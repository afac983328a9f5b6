// Tensor-core (tcgen05) version of the fused weather-MLP step: [standardise +] forward + cross-entropy + backward
// (+ DDP exchange + Adam).  Replaces /root/reference/jobs/train_lightning_ddp.py:57-71,:88,:136 (and, with the
// raw-column loader, /root/reference/jobs/preprocess.py:38-41) like mlp_grid_kernel, with the two F x 64 contractions
// on the 5th-generation tensor cores and as little per-row CUDA-core work as the arithmetic allows:
//
//   GEMM1   z[128 x 64]    = D[128 x K1] . W1s[64 x K1]^T           M=128 N=64  K1 = ceil((F+1)/8)*8   A from TMEM, B K-major SW128
//   GEMM2   T[64 x 2 K1]  += Gate[128 x 64]^T . [U_hi | U_lo]       M=64  N=2K1 K = 128 rows           MN-major, SW128_B32
//
//   D     = [d_0 .. d_{F-1}, 1, 0..]      d_c = x_c - float(mean_c): the raw column, centred only.  The scale 1/std and
//           the fp32 remainder of the mean are folded into the weights once per launch:
//           W1s[n] = [W1[n][c] / std_c .., b1[n] - sum_c W1[n][c] (mean_c - float(mean_c)) / std_c]   (computed in fp64)
//           so standardisation ((x - mean) / std, preprocess.py:38-41) costs ONE subtraction per cell on the CUDA cores.
//   Gate  = 1.0 where the unit is active (z > 0 and kept by dropout), else 0.0 -- exact in tf32, so GEMM2 needs no
//           hi/lo split of its A operand and the same Gate tile serves the hi and the lo half of U in ONE pass.
//   U     = delta * D with delta = (dl0 - dl1) / 2: for two classes the logit gradients are dl1 = -dl0 (softmax - onehot
//           sums to zero; their fp32 values differ by a rounding of 1/batch at most), hence with ds = 1/(1-p)
//               dh[r][j]  = gate[r][j] * delta[r] * ds (W2[0][j] - W2[1][j])
//               dW1[j][c] = ds (W2[0][j] - W2[1][j]) / std_c * (T[j][c] - (mean_c - float(mean_c)) T[j][F]),   db1[j] = (..) T[j][F]
//           applied once per CTA when T leaves tensor memory.  T accumulates in TMEM over ALL tiles of the CTA.
//
// fp32 accuracy on tf32 tensor cores: x = hi + lo (hi = x rounded to 11 significant bits, lo exact); GEMM1 sums
// hi*lo + lo*hi + hi*hi (~2^-21 relative); GEMM2's A operand is exact, so hi + lo of U alone gives the same accuracy.
// dW2 / db2 / loss are row-local sums: dW2 lives in per-thread TMEM columns (packed fp32x2 FMAs, two units each), db2 / loss
// in registers.
//
// Roles (480 threads, one CTA per SM, persistent over 128-row tiles; tile i belongs to worker group i & 1):
//   warps 0-3 / 4-7  two row-worker groups: thread r <-> tile row r <-> TMEM lane r (row-local work, no shuffles):
//                    everything from the GEMM1 accumulator on (ReLU/dropout/logits/softmax/dW2/Gate/U)
//   warps 8-11       operand builders, every tile: centre + hi/lo split of the landed row (GEMM1 A), Philox words
//   warp 12          TMA producer (cp.async.bulk of rows or F column segments + labels [+ keep bits], 2-3 stage ring)
//   warps 13 / 14    MMA issuers for GEMM1 / GEMM2 (whole warp runs the loop, elect.sync picks the issuing lane)
// Tail: CTA slabs are summed by a two-level ticket tree (12 CTAs per group, then the groups) in fixed order -- the
// result is bit-reproducible -- then the cross-rank exchange (csrc/peer.cuh) and Adam run in the last CTA.
#pragma once
#include <math.h>
#include <stddef.h>
#include <string.h>

#include "mlp_kernels.cuh"
#include "tc_common.cuh"

namespace wxtcstep {
using namespace wxtc;
using wxk::AdamScalars;
using wxk::GridArgs;

constexpr int TILE = 128;
constexpr int NGROUPS = 2;
constexpr int WORKERS = NGROUPS * TILE;          // 256 row-worker threads
constexpr int HELPERS = TILE;                    // 128 operand-builder threads (one per tile row)
constexpr int TC_THREADS = WORKERS + HELPERS + 96;   // + producer warp + two MMA warps
constexpr int HELPER_WARP0 = WORKERS / 32;
constexpr int PRODUCER_WARP = (WORKERS + HELPERS) / 32, MMA1_WARP = PRODUCER_WARP + 1, MMA2_WARP = PRODUCER_WARP + 2;
constexpr int MAX_STAGES = 6;
constexpr int MAXF_TC = 31;                      // F feature columns + the ones column fit the 32-column operand tile
constexpr int TMEM_COLS = 512;                   // D1 x3 [0,192) | T [192,256) | dW2 sums of group g [256+64g, 320+64g) |
                                                 // GEMM1 A operand (D hi | lo) of tile parity b [384+64b, 448+64b)
constexpr int ND1 = 3;                           // GEMM1 accumulators in flight (tile i uses D1[i % 3])
constexpr int NPW = 3;                           // Philox word buffers in flight (tile i uses pw[i % 3])
constexpr int TM_T = 192, TM_ACC = 256, TM_XA = 384;
constexpr int GROUP_CTAS = 12;                   // fan-in of the slab reduction tree
constexpr int TICKET_GROUP0 = 16, TICKET_TOP = 40;   // words of the ticket block (common.cuh: 64 words, 1/2/8 in use)
constexpr int SMEM_LIMIT = 232448 - 1024;        // 227 KB per CTA on sm_100, minus the static block (a flag, padded to the 1024-byte alignment of the dynamic part)

enum { DROP_OFF = WX_DROPOUT_OFF, DROP_MASK = WX_DROPOUT_MASK, DROP_PHILOX = WX_DROPOUT_PHILOX };

struct TcSmem {
    // operand tiles (all 1024-byte aligned)
    float w_hi[64 * 32], w_lo[64 * 32];          // W1s  [n][32]     K-major SW128         (GEMM1 B)
                                                 // (GEMM1's A operand, D hi/lo, lives in TMEM: thread r writes lane r)
    float gate[NGROUPS][TILE * 64];              // Gate [k=row][m=64]  MN-major SW128_B32 (GEMM2 A); reused by the final reductions
    float u[NGROUPS][TILE * 64];                 // U    [k=row][n: hi | lo] MN-major SW128_B32 (GEMM2 B); u[0]/u[1] hold the slab / bucket at the end
    float2 w2p[64];                              // ds * (W2[0][j], W2[1][j])
    __align__(16) float wdl[64];                 // ds * (W2[1][j] - W2[0][j]): the logit DIFFERENCE is all a 2-class softmax needs
    float b2[2];
    __align__(16) float mean_hi[32];             // raw loader: float(mean_c)
    float col_scale[32], col_shift[32];          // epilogue: 1/std_c and (mean_c - float(mean_c)) as floats
    float red_small[WORKERS / 32][4];            // db2_0, db2_1, loss per worker warp
    __align__(16) uint32_t pw[NPW][TILE][8];     // Philox words of the tiles in flight (helper -> workers), ring by tile
    uint64_t land_full[MAX_STAGES], land_empty[MAX_STAGES];
    uint64_t xa_ready[2], a2_ready[NGROUPS], a2_free[NGROUPS], d1_full[ND1], d1_free[ND1], pw_ready[NPW], pw_free[NPW];
    uint32_t tmem_base;
    int err;
    int role;                                    // 0 = not last, 1 = last of its group, 2 = last of the grid
    AdamScalars adam;
    __align__(128) unsigned char landing[128];   // ring: stages x { x[TILE*F] fp32 | y[TILE] i64 | keep[TILE] u64 } (dynamic size)
};

__device__ __forceinline__ void st4(void* base, uint32_t off, float a, float b, float c, float d) {
    *reinterpret_cast<float4*>(static_cast<unsigned char*>(base) + off) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    float a[16], b[16];
    tmem_ld16(taddr, a);
    tmem_ld16(taddr + 16, b);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        v[i] = a[i];
        v[16 + i] = b[i];
    }
}

struct TcGeom {
    int stages;          // landing ring depth (2 or 3)
    int x_bytes;         // TILE * F * 4 rounded up to 128
    int stage_bytes;     // x_bytes + 2 * TILE * 8
};

// RAW : raw-column loader (column segments [F][128] in the landing buffer, centred on the way) vs dense row-major rows.
// DROP: dropout mode -- compile-time so the per-unit keep test is one multiply (Philox digits) or one bit test (mask).
// FT  : the feature count when it is known at compile time (column loops unroll, operand offsets fold), 0 = use F_rt.
template <bool RAW, int DROP, int FT>
__global__ void __launch_bounds__(TC_THREADS, 1) tc_step_kernel(const GridArgs a, const int F_rt, const TcGeom geo,
                                                                int* __restrict__ err_out) {
    const int F = FT ? FT : F_rt;
    constexpr int UNR = FT ? 4 : 1;               // unroll factor of the column-group loops
    extern __shared__ __align__(1024) unsigned char raw[];     // 1024-byte alignment: the swizzle atoms need it
    TcSmem& s = *reinterpret_cast<TcSmem*>(raw);               // direct cast keeps the shared address space (LDS/STS)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int P = 64 * F + 194;
    const int OFF_B1 = 64 * F, OFF_W2 = 64 * F + 64, OFF_B2 = 64 * F + 192;
    const int slab_stride = (P + 1 + 3) & ~3;
    const int K1 = (F + 1 + 7) & ~7;              // GEMM1 K extent: features | 1 at column F | zeros
    const int N2 = 2 * K1;                        // GEMM2 N extent: [U_hi | U_lo]
    const int nfull = F >> 3;                     // column groups of 8 that hold features only
    const int64_t ntiles = (a.batch + TILE - 1) / TILE;
    // tiles of this CTA: >= 1 by construction of the grid, far below 2^31 -- 32-bit loop counters keep the ring
    // arithmetic (i % 3, ...) off the 64-bit division path
    const int my_tiles = int((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x);
    const float ds = DROP != DROP_OFF ? 1.25f : 1.0f;           // 1/(1-0.2), train_lightning_ddp.py:60

#ifdef WX_TC_TIMERS
#define WX_TC_STAMP(k)                                                   \
    do {                                                                 \
        if (a.dbg != nullptr && tid == 0) {                              \
            long long t__;                                               \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));      \
            a.dbg[blockIdx.x * 8 + (k)] = t__;                           \
        }                                                                \
    } while (0)
#else
#define WX_TC_STAMP(k) do { } while (0)
#endif
#ifdef WX_TC_TIMERS
    // per-role phase timers: cycles accumulated by one thread of each role, written to dbg[148*8 + (cta*4 + role)*8 + k]
    long long ph__[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long mark__ = clock64();
#define WX_TC_PH(k)                          \
    do {                                     \
        const long long now__ = clock64();   \
        ph__[k] += now__ - mark__;           \
        mark__ = now__;                      \
    } while (0)
#define WX_TC_PH_FLUSH(role)                                                                        \
    do {                                                                                            \
        if (a.dbg != nullptr)                                                                       \
            for (int k__ = 0; k__ < 8; ++k__) a.dbg[148 * 8 + (blockIdx.x * 5 + (role)) * 8 + k__] = ph__[k__]; \
    } while (0)
#else
#define WX_TC_PH(k) do { } while (0)
#define WX_TC_PH_FLUSH(role) do { } while (0)
#endif
    WX_TC_STAMP(0);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the next launch may begin its own prologue
    if (tid == 0 && (smem_u32(raw) & 1023u) != 0) atomicExch(err_out, 99);   // dynamic smem base not 1024-aligned
    if (tid == 0) {
        for (int i = 0; i < MAX_STAGES; ++i) {
            mbar_init(&s.land_full[i], 32);
            mbar_init(&s.land_empty[i], TILE);
        }
        for (int g = 0; g < NGROUPS; ++g) {
            mbar_init(&s.xa_ready[g], TILE);
            mbar_init(&s.a2_ready[g], TILE);
            mbar_init(&s.a2_free[g], 1);
        }
        for (int g = 0; g < NPW; ++g) {
            mbar_init(&s.pw_ready[g], TILE);
            mbar_init(&s.pw_free[g], TILE);
        }
        for (int g = 0; g < ND1; ++g) {
            mbar_init(&s.d1_full[g], 1);
            mbar_init(&s.d1_free[g], TILE);
        }
        fence_mbar_init();
        s.err = 0;
    }
    if (warp == 0) tmem_alloc(&s.tmem_base, TMEM_COLS);
    __syncthreads();                                              // the barriers exist for every role from here on
    // Everything above touches only this CTA's own resources and may overlap the tail of the previous launch in the
    // stream (programmatic dependent launch); inputs, parameters, statistics, step counter and scratch are read below.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    // The producer warp goes straight to its loop: the first tiles are on their way from HBM while the other 14 warps
    // stage the parameters and build the W1s operand (they meet on named barrier 2; everyone meets again at the tail).
    constexpr int NSTAGE_THREADS = TC_THREADS - 32;
    const int stid = tid < PRODUCER_WARP * 32 ? tid : tid - 32;   // index among the staging threads
    float* sp = s.u[0];                                           // [P] parameters
    double* sst = reinterpret_cast<double*>(s.u[0] + ((P + 3) & ~3) + 4);   // [F][4] statistics (8-byte aligned)
    if (warp != PRODUCER_WARP) {
    // parameters and statistics: one coalesced pass into shared memory (U[0] is free until the first tile ends)
    for (int i = stid; i < P; i += NSTAGE_THREADS) sp[i] = a.params[i];
    if (RAW)
        for (int i = stid; i < 4 * F; i += NSTAGE_THREADS) sst[i] = a.stats[i];
    if (tid == TC_THREADS - 1 && a.tail == wxk::TAIL_ADAM) s.adam = wxk::adam_scalars(a.desc, a.state[0] + 1);
    asm volatile("bar.sync 2, %0;" ::"n"(NSTAGE_THREADS) : "memory");
    // W1s operand (hi/lo, K-major SW128), one thread per hidden unit; fp64 so that folding 1/std and the mean remainder
    // into the weights costs no accuracy (the products are rounded to fp32 once, like W1 itself)
    if (tid < 64) {
        const int n = tid;
        double shift = 0.0;
#pragma unroll 1
        for (int c4 = 0; c4 < 8; ++c4) {
            float hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int c = c4 * 4 + e;
                float v = 0.0f;
                if (c < F) {
                    double w = double(sp[n * F + c]);
                    if (RAW) {
                        const double mean = sst[c * 4 + 1], rs = 1.0 / sst[c * 4 + 3];
                        w *= rs;
                        shift += w * (mean - double(float(mean)));
                    }
                    v = float(w);
                }
                split_tf32(v, hi[e], lo[e]);
            }
            const uint32_t off = sw128_offset(n, c4 * 4);
            st4(s.w_hi, off, hi[0], hi[1], hi[2], hi[3]);
            st4(s.w_lo, off, lo[0], lo[1], lo[2], lo[3]);
        }
        float bh, bl;
        split_tf32(float(double(sp[OFF_B1 + n]) - shift), bh, bl);
        const uint32_t off = sw128_offset(n, F);
        *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(s.w_hi) + off) = bh;
        *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(s.w_lo) + off) = bl;
        s.w2p[n] = make_float2(ds * sp[OFF_W2 + n], ds * sp[OFF_W2 + 64 + n]);
        s.wdl[n] = ds * (sp[OFF_W2 + 64 + n] - sp[OFF_W2 + n]);
    } else if (tid < 96) {
        const int c = tid - 64;
        float mh = 0.0f, sc = 1.0f, sh = 0.0f;
        if (RAW && c < F) {
            const double mean = sst[c * 4 + 1];
            mh = float(mean);
            sc = float(1.0 / sst[c * 4 + 3]);
            sh = float(mean - double(mh));
        }
        s.mean_hi[c] = mh;
        s.col_scale[c] = sc;
        s.col_shift[c] = sh;
        if (c < 2) s.b2[c] = sp[OFF_B2 + c];
    }
    fence_proxy_async();
    tc_fence_before();
    asm volatile("bar.sync 2, %0;" ::"n"(NSTAGE_THREADS) : "memory");
    tc_fence_after();
    }
    WX_TC_STAMP(1);
    const uint32_t tm = s.tmem_base;
    const uint64_t step_word = (a.state != nullptr) ? uint64_t(a.state[0]) : a.step;

    auto tile_rows = [&](int64_t t) -> int {
        const int64_t left = a.batch - t * TILE;
        return left < TILE ? int(left) : TILE;
    };
    auto land_x = [&](int st) -> float* { return reinterpret_cast<float*>(s.landing + size_t(st) * geo.stage_bytes); };
    auto land_y = [&](int st) -> long long* {
        return reinterpret_cast<long long*>(s.landing + size_t(st) * geo.stage_bytes + geo.x_bytes);
    };
    auto land_m = [&](int st) -> unsigned long long* {
        return reinterpret_cast<unsigned long long*>(s.landing + size_t(st) * geo.stage_bytes + geo.x_bytes + TILE * 8);
    };

    if (warp == PRODUCER_WARP) {
        // ================= TMA producer =================
        const int S = geo.stages;
        int st = 0;
        uint32_t round = 0;                      // how many times the ring has wrapped
        for (int i = 0; i < my_tiles; ++i) {
            const int64_t tile = int64_t(blockIdx.x) + int64_t(i) * gridDim.x;
            WX_TC_PH(7);
            if (round > 0 && !mbar_wait(&s.land_empty[st], (round - 1) & 1u)) s.err = 2;
            WX_TC_PH(0);
            const int nr = tile_rows(tile);
            const int64_t row = a.row0 + tile * TILE;
            float* lx = land_x(st);
            if (nr == TILE && RAW) {
                // full tile of raw columns: F column segments of 512 B + labels [+ keep bits].  As bulk-TMA copies these
                // cost ~115 cycles EACH to issue (22 per tile: the producer became the pipeline's bottleneck); as 16-byte
                // cp.async they are one instruction per 512 B for the whole warp: lane l moves rows 4l .. 4l+3 of every
                // column, and its completion arrive on the stage barrier follows its own copies.
#pragma unroll UNR
                for (int c = 0; c < F; ++c)
                    cp_async16(lx + c * TILE + 4 * lane, static_cast<const float*>(a.cols.col[c]) + row + 4 * lane);
                cp_async16(land_y(st) + 2 * lane, a.y + row + 2 * lane);
                cp_async16(land_y(st) + 64 + 2 * lane, a.y + row + 64 + 2 * lane);
                if (DROP == DROP_MASK) {
                    cp_async16(land_m(st) + 2 * lane, a.keep_mask + tile * TILE + 2 * lane);
                    cp_async16(land_m(st) + 64 + 2 * lane, a.keep_mask + tile * TILE + 64 + 2 * lane);
                }
                cp_async_mbar_arrive(&s.land_full[st]);
            } else if (nr == TILE) {    // full tile of row-major rows: one contiguous bulk copy (+ labels [+ keep bits])
                const uint32_t by = TILE * 8;
                if (lane == 0) {
                    mbar_arrive_expect_tx(&s.land_full[st], uint32_t(TILE) * F * 4 + by + (DROP == DROP_MASK ? by : 0u));
                    bulk_g2s(lx, a.x + row * F, uint32_t(TILE) * F * 4, &s.land_full[st]);
                    bulk_g2s(land_y(st), a.y + row, by, &s.land_full[st]);
                    if (DROP == DROP_MASK) bulk_g2s(land_m(st), a.keep_mask + tile * TILE, by, &s.land_full[st]);
                }
                __syncwarp();
                if (lane != 0) mbar_arrive(&s.land_full[st]);     // the barrier counts the whole warp (see the cp.async path)
            } else {
                // ragged last tile: plain loads by the whole warp; rows past the end are filled with values that centre
                // to exactly zero (they must not put NaN into the operand tiles), then a normal arrive
                if (!RAW) {
                    const float* gx = a.x + row * F;
                    for (int e = lane; e < TILE * F; e += 32) lx[e] = e < nr * F ? gx[e] : 0.0f;
                } else {
                    for (int c = 0; c < F; ++c)
                        for (int e = lane; e < TILE; e += 32)
                            lx[c * TILE + e] = e < nr ? static_cast<const float*>(a.cols.col[c])[row + e]
                                                      : float(a.stats[c * 4 + 1]);   // centres to exactly zero (= s.mean_hi[c], which
                                                                                     // the staging warps may not have written yet)
                }
                for (int e = lane; e < TILE; e += 32) {
                    land_y(st)[e] = e < nr ? a.y[row + e] : 0;
                    if (DROP == DROP_MASK) land_m(st)[e] = e < nr ? a.keep_mask[tile * TILE + e] : 0ull;
                }
                __threadfence_block();
                mbar_arrive(&s.land_full[st]);
            }
            WX_TC_PH(1);
            if (++st == S) {
                st = 0;
                ++round;
            }
        }
        if (lane == 0) WX_TC_PH_FLUSH(4);
    } else if (warp == MMA1_WARP) {
        // ================= MMA issuer 1: GEMM1 of every tile =================
        // The WHOLE warp runs this code and one elected lane issues each MMA: operands stay warp-uniform (uniform
        // registers) -- 3x faster issue than a lone thread.  GEMM1 and GEMM2 are issued by different warps so that a
        // ready GEMM1 (9 MMAs, a row-worker group is waiting for it) is not queued behind the 16 MMAs of a GEMM2.
        const uint32_t id1 = make_idesc_tf32(128, 64, false, false);
        const uint64_t wk_d[2] = {make_smem_desc(smem_u32(s.w_hi), 16, 1024), make_smem_desc(smem_u32(s.w_lo), 16, 1024)};
        const int k1steps = K1 / 8;
        WX_TC_PH(7);
        for (int i = 0; i < my_tiles; ++i) {
            const int q = i % ND1;
            if (!mbar_wait_hot(&s.xa_ready[i & 1], uint32_t((i >> 1) & 1))) s.err = 3;
            // D1[q] is free once the row workers of tile i - 3 have read it for the last time
            if (i >= ND1 && !mbar_wait_hot(&s.d1_free[q], uint32_t((i / ND1 - 1) & 1))) s.err = 13;
            tc_fence_after();
            WX_TC_PH(0);
            const uint32_t dst = tm + uint32_t(q) * 64, xa = tm + TM_XA + uint32_t(i & 1) * 64;
            bool acc = false;                    // small terms first: hi*lo, lo*hi, then hi*hi
#pragma unroll
            for (int term = 0; term < 3; ++term) {
                const uint32_t at = xa + (term == 1 ? 32u : 0u);     // A from TMEM: hi at columns [0,32), lo at [32,64)
                const uint64_t bd = wk_d[term == 0 ? 1 : 0];
#pragma unroll UNR
                for (int k = 0; k < k1steps; ++k) {
                    umma_tf32_ts_warp(dst, at + uint32_t(k * 8), bd + uint64_t(k * 2), id1, acc);  // 8 columns / 32 B per k-step
                    acc = true;
                }
            }
            umma_commit_warp(&s.d1_full[q]);
            WX_TC_PH(1);
        }
        if (lane == 0) WX_TC_PH_FLUSH(3);
    } else if (warp == MMA2_WARP) {
        // ================= MMA issuer 2: GEMM2 of every tile, as its Gate / U tiles become ready =================
        const uint32_t id2 = make_idesc_tf32(64, N2, true, true);
        // group 1's tiles sit 32 KB after group 0's: + 2048 in the start-address field (16-byte units)
        const uint64_t gate_d0 = make_smem_desc(smem_u32(s.gate[0]), 512, 1024, LAYOUT_SW128_B32);
        const uint64_t u_d0 = make_smem_desc(smem_u32(s.u[0]), 512, 1024, LAYOUT_SW128_B32);
        constexpr uint64_t GROUP_STEP = sizeof(float) * TILE * 64 / 16;
        for (int i = 0; i < my_tiles; ++i) {
            const int g = i & 1;
            if (!mbar_wait_hot(&s.a2_ready[g], uint32_t((i >> 1) & 1))) s.err = 4;
            tc_fence_after();
            const uint64_t gd = gate_d0 + uint64_t(g) * GROUP_STEP, ud = u_d0 + uint64_t(g) * GROUP_STEP;
#pragma unroll
            for (int k = 0; k < TILE / 8; ++k)
                umma_tf32_warp(tm + TM_T, gd + uint64_t(k * 128), ud + uint64_t(k * 128), id2, !(i == 0 && k == 0));   // +2048 B
            umma_commit_warp(&s.a2_free[g]);
        }
    } else if (warp >= HELPER_WARP0) {
        // ================= operand builders (4 warps, thread r <-> tile row r, EVERY tile) =================
        // Centre the landed row, split it into the K-major hi/lo operand of GEMM1, and draw the row's dropout words --
        // everything a tile needs before its GEMM1, so the row workers start at the accumulator.
        const int r = tid - HELPER_WARP0 * 32;
        const uint2 key = make_uint2(uint32_t(a.seed), uint32_t(a.seed >> 32));
        const bool impute = a.impute == WX_IMPUTE_MEAN;
        constexpr int cstride = RAW ? TILE : 1;
        const uint32_t lane_addr = uint32_t(((warp & 3) * 32)) << 16;   // helper thread r owns TMEM lane r
        int st = 0;
        uint32_t land_par = 0;                                    // landing ring position / phase of tile i
        for (int i = 0; i < my_tiles; ++i) {
            const int64_t tile = int64_t(blockIdx.x) + int64_t(i) * gridDim.x;
            WX_TC_PH(7);
            // ---- dropout words of this row (independent of everything else: drawn first, handed over when the ring
            //      slot has been read by the workers of tile i - 3) ----
            if (DROP == DROP_PHILOX) {
                uint32_t pw[8];
                wxk::philox_keep_words(uint64_t(tile * TILE + r), key, step_word, pw);
                const int q = i % NPW;
                if (i >= NPW && !mbar_wait(&s.pw_free[q], uint32_t((i / NPW - 1) & 1))) s.err = 12;
                *reinterpret_cast<uint4*>(&s.pw[q][r][0]) = make_uint4(pw[0], pw[1], pw[2], pw[3]);
                *reinterpret_cast<uint4*>(&s.pw[q][r][4]) = make_uint4(pw[4], pw[5], pw[6], pw[7]);
                mbar_arrive(&s.pw_ready[q]);
            }
            WX_TC_PH(0);
            if (!mbar_wait(&s.land_full[st], land_par)) s.err = 5;
            WX_TC_PH(1);
            // the A operand buffer of this tile parity is free once GEMM1 of tile i - 2 has completed
            if (i >= 2 && !mbar_wait(&s.d1_full[(i - 2) % ND1], uint32_t(((i - 2) / ND1) & 1))) s.err = 10;
            tc_fence_after();
            WX_TC_PH(2);
            float* lx = land_x(st) + (RAW ? r : r * F);           // element c of this row at lx[c * cstride]
            const uint32_t xa = tm + TM_XA + uint32_t(i & 1) * 64 + lane_addr;

            // ---- D row -> hi/lo A operand of GEMM1 in TMEM (lane = row, column = feature).  Raw columns are centred here
            //      (one subtraction per cell; nulls become 0 = the column mean under WX_IMPUTE_MEAN) and written back to
            //      the landing buffer for their second use in U ----
            // all loads first, all stores last: the landing ring and the constants live in the same shared array, so a
            // store between two loads would serialise them
            float dh[32], dlo[32], xv[32], mh[32];
#pragma unroll
            for (int c = 0; c < 32; ++c) xv[c] = (c < K1 && c < F) ? lx[c * cstride] : 0.0f;
            if (RAW) {
#pragma unroll
                for (int c4 = 0; c4 < 8; ++c4) {
                    if (c4 * 4 < F) {
                        const float4 m = *reinterpret_cast<const float4*>(&s.mean_hi[c4 * 4]);
                        mh[c4 * 4] = m.x; mh[c4 * 4 + 1] = m.y; mh[c4 * 4 + 2] = m.z; mh[c4 * 4 + 3] = m.w;
                    } else {
                        mh[c4 * 4] = mh[c4 * 4 + 1] = mh[c4 * 4 + 2] = mh[c4 * 4 + 3] = 0.0f;
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < 32; ++c) {
                float d = 0.0f;
                if (c < K1) {                                     // (compile-time for the specialised builds)
                    if (c < F) {
                        d = xv[c];
                        if (RAW) {
                            d = xv[c] - mh[c];
                            if (impute) d = (xv[c] != xv[c]) ? 0.0f : d;
                            xv[c] = d;
                        }
                    } else if (c == F) {
                        d = 1.0f;
                    }
                }
                split_tf32(d, dh[c], dlo[c]);
            }
            if (RAW) {
#pragma unroll
                for (int c = 0; c < 32; ++c)
                    if (c < K1 && c < F) lx[c * cstride] = xv[c];
            }
#pragma unroll
            for (int c16 = 0; c16 < 2; ++c16) {
                if (c16 * 16 < K1) {
                    float h16[16], l16[16];
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        h16[e] = dh[c16 * 16 + e];
                        l16[e] = dlo[c16 * 16 + e];
                    }
                    tmem_st16_nc(xa + c16 * 16, h16);
                    tmem_st16_nc(xa + 32 + c16 * 16, l16);
                }
            }
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(&s.xa_ready[i & 1]);
            WX_TC_PH(3);
            if (++st == geo.stages) {
                st = 0;
                land_par ^= 1u;
            }
        }
        if (r == 0) WX_TC_PH_FLUSH(2);
    } else {
        // ================= row workers: group 0 = warps 0-3, group 1 = warps 4-7 =================
        const int grp = warp >> 2;
        const int r = tid & (TILE - 1);                           // tile row == TMEM lane
        const uint32_t lane_addr = uint32_t((warp & 3) * 32) << 16;
        const uint32_t acc_base = tm + TM_ACC + uint32_t(grp) * 64 + lane_addr;
        float gb2_0 = 0.0f, gb2_1 = 0.0f, loss_acc = 0.0f;
        const float b2_0 = s.b2[0], b2_1 = s.b2[1];
        constexpr int cstride = RAW ? TILE : 1;
        {   // zero this thread's dW2 sums in TMEM
            float zero[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) zero[q] = 0.0f;
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_st16(acc_base + c * 16, zero);
            tmem_st_wait();
        }

        int it = 0;                                               // tiles this group has processed
        int st = grp % geo.stages;                                // landing ring position of tile i (advances by 2)
        for (int i = grp; i < my_tiles; i += NGROUPS, ++it) {
            const int64_t tile = int64_t(blockIdx.x) + int64_t(i) * gridDim.x;
            const uint32_t par = uint32_t(it & 1);                // parity of this group's per-tile barriers
            float* lx = land_x(st) + (RAW ? r : r * F);           // element c of this (already centred) row at lx[c * cstride]
            // ---- dropout words of this row, made by the helper warps while this group finished its previous tile ----
            uint32_t pw[8];
            WX_TC_PH(7);
            if (DROP == DROP_PHILOX) {
                const int q = i % NPW;
                if (!mbar_wait(&s.pw_ready[q], uint32_t((i / NPW) & 1))) s.err = 11;
                const uint4 p0 = *reinterpret_cast<const uint4*>(&s.pw[q][r][0]);
                const uint4 p1 = *reinterpret_cast<const uint4*>(&s.pw[q][r][4]);
                pw[0] = p0.x; pw[1] = p0.y; pw[2] = p0.z; pw[3] = p0.w;
                pw[4] = p1.x; pw[5] = p1.y; pw[6] = p1.z; pw[7] = p1.w;
                mbar_arrive(&s.pw_free[q]);
            }
            WX_TC_PH(0);
            WX_TC_PH(1);
            const uint32_t d1 = tm + uint32_t(i % ND1) * 64 + lane_addr;
            if (!mbar_wait(&s.d1_full[i % ND1], uint32_t((i / ND1) & 1))) s.err = 6;   // GEMM1 done => stage filled and centred
            tc_fence_after();
            WX_TC_PH(2);
            uint64_t keep = ~0ull;
            if (DROP == DROP_MASK) keep = land_m(st)[r];
            const int yv = int(land_y(st)[r]);
            const bool valid = r < tile_rows(tile);

            // ---- phase A: a = relu(z) where kept (dropout scale folded into W2 and the dlogits), logits, Gate tile.
            //      The TMEM load of the next 16 units is in flight while this chunk is processed. ----
            // Only lg1 - lg0 enters a two-class log-softmax, so one FMA per unit against ds (W2[1][j] - W2[0][j]) replaces the
            // two logit FMAs; the 16 weights of the next chunk are requested before this chunk is processed.
            // four independent accumulation chains (units = 0..3 mod 4), two per packed fp32x2 FMA
            float2 dq2[2] = {make_float2(0.0f, 0.0f), make_float2(0.0f, 0.0f)};
            const float4* __restrict__ wdl4 = reinterpret_cast<const float4*>(s.wdl);   // read-only here
            float zbuf[2][16];
            float4 wbuf[2][4];
            tmem_ld16_nc(d1, zbuf[0]);
#pragma unroll
            for (int k = 0; k < 4; ++k) wbuf[0][k] = wdl4[k];
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) {
                float (&z)[16] = zbuf[cc & 1];
                tmem_ld_wait_dep(z);
                if (cc < 3) {
                    tmem_ld16_nc(d1 + (cc + 1) * 16, zbuf[(cc + 1) & 1]);
#pragma unroll
                    for (int k = 0; k < 4; ++k) wbuf[(cc + 1) & 1][k] = wdl4[(cc + 1) * 4 + k];
                }
                const float4 (&w4)[4] = wbuf[cc & 1];
                const float wv[16] = {w4[0].x, w4[0].y, w4[0].z, w4[0].w, w4[1].x, w4[1].y, w4[1].z, w4[1].w,
                                      w4[2].x, w4[2].y, w4[2].z, w4[2].w, w4[3].x, w4[3].y, w4[3].z, w4[3].w};
                uint32_t s0 = 0, s1 = 0;
                if (DROP == DROP_PHILOX) {
                    s0 = pw[2 * cc];
                    s1 = pw[2 * cc + 1];
                }
                const uint32_t kb = uint32_t(keep >> (cc * 16)) & 0xffffu;
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    bool kp = true;
                    if (DROP == DROP_PHILOX) kp = wxk::keep_digit(q < 8 ? s0 : s1);
                    if (DROP == DROP_MASK) kp = (kb >> q) & 1u;
                    const bool on = kp && !(z[q] <= 0.0f);                 // relu as torch's threshold: NaN passes
                    z[q] = on ? z[q] : 0.0f;
                    if (q & 1)
                        dq2[(q >> 1) & 1] = __ffma2_rn(make_float2(z[q - 1], z[q]), make_float2(wv[q - 1], wv[q]), dq2[(q >> 1) & 1]);
                }
                tmem_st16_nc(d1 + cc * 16, z);                               // a replaces z in place: phase B reads it back
            }
            WX_TC_PH(3);
            const float lg0 = 0.0f;                               // log-softmax is shift-invariant: (0, lg1 - lg0)
            const float lg1 = ((dq2[0].x + dq2[0].y) + (dq2[1].x + dq2[1].y)) + (b2_1 - b2_0);
            // ---- softmax / loss / dlogits, ATen's operation order ----
            const float mx = fmaxf(lg0, lg1);
            const float e0 = lg0 - mx, e1 = lg1 - mx;
            const float lsum = logf(expf(e0) + expf(e1));
            const float lp0 = e0 - lsum, lp1 = e1 - lsum;
            float dl0 = 0.0f, dl1 = 0.0f;
            if (valid) {
                loss_acc += -(yv ? lp1 : lp0);
                dl0 = (expf(lp0) - (yv == 0 ? 1.0f : 0.0f)) * a.scale;
                dl1 = (expf(lp1) - (yv == 1 ? 1.0f : 0.0f)) * a.scale;
                gb2_0 += dl0;
                gb2_1 += dl1;
            }
            const float delta = 0.5f * (dl0 - dl1);              // dl1 = -dl0 up to rounding (two classes); ds rides in w2p
            WX_TC_PH(4);
            // ---- phase B: dW2[0][j] sum, g[j] += a_j * ds * delta, in this thread's TMEM columns (dW2[1] = -dW2[0]), and the
            //      Gate tile (a_j != 0 <=> the unit is active).  Gate / U of this group are free once GEMM2 of its previous
            //      tile has drained: it was issued when that tile ended, a whole phase A ago. ----
            tmem_st_wait();
            if (it > 0 && !mbar_wait(&s.a2_free[grp], par ^ 1u)) s.err = 7;
            {
                const float dd = ds * delta;
                float abuf[2][16], gbuf[2][16];
                tmem_ld16_nc(d1, abuf[0]);
                tmem_ld16_nc(acc_base, gbuf[0]);
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) {
                    float (&av)[16] = abuf[cc & 1];
                    float (&g)[16] = gbuf[cc & 1];
                    tmem_ld_wait_dep(av);
                    tmem_dep(g);
                    if (cc < 3) {
                        tmem_ld16_nc(d1 + (cc + 1) * 16, abuf[(cc + 1) & 1]);
                        tmem_ld16_nc(acc_base + (cc + 1) * 16, gbuf[(cc + 1) & 1]);
                    }
#pragma unroll
                    for (int q = 0; q < 16; q += 2) {
                        const float2 t = __ffma2_rn(make_float2(av[q], av[q + 1]), make_float2(dd, dd), make_float2(g[q], g[q + 1]));
                        g[q] = t.x;
                        g[q + 1] = t.y;
                    }
                    tmem_st16_nc(acc_base + cc * 16, g);
                    float gt[16];
#pragma unroll
                    for (int q = 0; q < 16; ++q) gt[q] = (av[q] == 0.0f) ? 0.0f : 1.0f;    // a >= 0 or NaN (NaN counts as active)
#pragma unroll
                    for (int g8 = 0; g8 < 2; ++g8) {
                        const uint32_t off = mn32_offset(cc * 16 + g8 * 8, r, 512, 1024);
                        st4(s.gate[grp], off, gt[g8 * 8], gt[g8 * 8 + 1], gt[g8 * 8 + 2], gt[g8 * 8 + 3]);
                        st4(s.gate[grp], off + 16, gt[g8 * 8 + 4], gt[g8 * 8 + 5], gt[g8 * 8 + 6], gt[g8 * 8 + 7]);
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(&s.d1_free[i % ND1]);                     // last read of this tile's accumulator is done
            WX_TC_PH(5);
            // ---- U row = delta * D (MN-major, hi at columns [0,K1), lo at [K1, 2 K1)) ----
            {
                unsigned char* ub = reinterpret_cast<unsigned char*>(s.u[grp]);
                // all loads of the centred row first (a store between two loads of the same shared array serialises them)
                float dv[32];
#pragma unroll
                for (int c = 0; c < 32; ++c) dv[c] = (c < K1 && c < F) ? lx[c * cstride] : 0.0f;
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    if (c8 * 8 < K1) {
                        float hi[8], lo[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const int c = c8 * 8 + e;
                            const float v = c < F ? delta * dv[c] : (c == F ? delta : 0.0f);
                            split_tf32(v, hi[e], lo[e]);
                        }
                        const uint32_t oh = mn32_offset(c8 * 8, r, 512, 1024), ol = mn32_offset(K1 + c8 * 8, r, 512, 1024);
                        st4(ub, oh, hi[0], hi[1], hi[2], hi[3]);
                        st4(ub, oh + 16, hi[4], hi[5], hi[6], hi[7]);
                        st4(ub, ol, lo[0], lo[1], lo[2], lo[3]);
                        st4(ub, ol + 16, lo[4], lo[5], lo[6], lo[7]);
                    }
                }
            }
            tmem_st_wait();
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(&s.a2_ready[grp]);
            mbar_arrive(&s.land_empty[st]);
            st += NGROUPS;
            if (st >= geo.stages) st -= geo.stages;
            WX_TC_PH(6);
        }
        if (r == 0) WX_TC_PH_FLUSH(grp);
        WX_TC_STAMP(2);
        // ---- all tiles done: wait for this group's last GEMM2, then (after both groups) read T and assemble the slab ----
        if (it > 0 && !mbar_wait(&s.a2_free[grp], uint32_t((it - 1) & 1))) s.err = 8;
        tc_fence_after();
        asm volatile("bar.sync 1, 256;" ::: "memory");
        float* slab = s.u[0];                                       // GEMM2 is done with both U buffers
        if (grp == 0) {
            // T[m][n] (M=64 accumulator: row m = 16*warp + lane, lanes 0-15 of each sub-partition), hi half + lo half
            float v0[32], v1[32];
            tmem_ld32(tm + TM_T + lane_addr, v0);
            tmem_ld32(tm + TM_T + lane_addr + 32, v1);
            tmem_ld_wait();
            if (lane < 16) {
                // through shared memory (U[1] is free): the column arithmetic below depends on the runtime K1 and F
                const int m = (warp & 3) * 16 + lane;
                float* trow = s.u[1] + m * 65;
#pragma unroll
                for (int q = 0; q < 32; ++q) {
                    trow[q] = v0[q];
                    trow[32 + q] = v1[q];
                }
                const float2 w = s.w2p[m];
                const float wd = w.x - w.y;                         // ds * (W2[0][m] - W2[1][m])
                const float tone = trow[F] + trow[K1 + F];          // T[m][F]: sum over rows of gate * delta
                slab[OFF_B1 + m] = wd * tone;
                for (int c = 0; c < F; ++c)
                    slab[m * F + c] = wd * s.col_scale[c] * ((trow[c] + trow[K1 + c]) - s.col_shift[c] * tone);
            }
        }
        // dW2[0][j]: per-thread sums (64 TMEM columns) -> ordered column sums through smem (the Gate tiles are free now).
        // Both groups store at once: row (grp, r) keeps its 64 values at column (c + 4 r) & 63, so the row-wise stores
        // and the column-wise reads are free of bank conflicts; thread (h, j) of group 0 then adds the 128 rows of
        // group h for column j in row order, and the two halves are added in group order.
        float* red = s.gate[0];                                       // [2][128 rows][64] floats = 64 KB = both Gate tiles
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            float v[16];
            tmem_ld16(acc_base + c * 16, v);
            tmem_ld_wait();
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4)
                *reinterpret_cast<float4*>(&red[(grp * TILE + r) * 64 + ((c * 16 + q4 * 4 + 4 * r) & 63)]) =
                    make_float4(v[q4 * 4], v[q4 * 4 + 1], v[q4 * 4 + 2], v[q4 * 4 + 3]);
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (grp == 0) {
            const int h = r >> 6, j = r & 63;
            float cs[4] = {0.0f, 0.0f, 0.0f, 0.0f};                     // four chains (rows k = q mod 4), added in a fixed order
#pragma unroll 8
            for (int k = 0; k < TILE; ++k) cs[k & 3] += red[(h * TILE + k) * 64 + ((j + 4 * k) & 63)];
            s.w_hi[r] = (cs[0] + cs[1]) + (cs[2] + cs[3]);              // W1s is dead: 128 partial column sums
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (grp == 0 && r < 64) {
            const float* part = s.w_hi;
            const float d = part[r] + part[64 + r];
            slab[OFF_W2 + r] = d;                                       // dW2[0][j]
            slab[OFF_W2 + 64 + r] = -d;                                 // dW2[1][j] = -dW2[0][j] (two classes)
        }
        // db2 / loss: warp sums in a fixed order
        const float w0 = warp_sum(gb2_0), w1 = warp_sum(gb2_1), w2 = warp_sum(loss_acc);
        if (lane == 0) {
            s.red_small[warp][0] = w0;
            s.red_small[warp][1] = w1;
            s.red_small[warp][2] = w2;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (tid < 3) {
            float t = 0.0f;
            for (int k = 0; k < WORKERS / 32; ++k) t += s.red_small[k][tid];
            slab[tid < 2 ? OFF_B2 + tid : P] = t;
        }
    }
    // ================= CTA slab -> global; two-level ticket tree sums the slabs in a fixed order =================
    tc_fence_before();
    __syncthreads();
    WX_TC_STAMP(3);
    if (warp == 0) tmem_dealloc(tm, TMEM_COLS);
    const float* slab = s.u[0];
    const unsigned nparts = gridDim.x;
    const unsigned ngroups = (nparts + GROUP_CTAS - 1) / GROUP_CTAS;
    const unsigned mygroup = blockIdx.x / GROUP_CTAS;
    float* mypart = a.part + size_t(blockIdx.x) * slab_stride;
    float* gpart = a.part + size_t(nparts) * slab_stride;             // [ngroups][slab_stride] group sums
    unsigned int* tick = a.ticket - 1;                                // word 0 of the ticket block
    for (int i = tid; i < slab_stride; i += TC_THREADS) mypart[i] = i < P + 1 ? slab[i] : 0.0f;
    if (tid == 0 && s.err != 0) atomicExch(err_out, s.err);
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned gsize = min(unsigned(GROUP_CTAS), nparts - mygroup * GROUP_CTAS);
        s.role = atomicAdd(&tick[TICKET_GROUP0 + mygroup], 1u) == gsize - 1 ? 1 : 0;
    }
    __syncthreads();
    WX_TC_STAMP(4);
    if (s.role == 0) return;
    __threadfence();
    const int ng4 = slab_stride / 4;
    {   // level 1: this CTA arrived last in its group -> sum the group's slabs in CTA order
        const unsigned b0 = mygroup * GROUP_CTAS, b1 = min(b0 + GROUP_CTAS, nparts);
        for (int g = tid; g < ng4; g += TC_THREADS) {
            const float4* src = reinterpret_cast<const float4*>(a.part) + g;
            float4 t[GROUP_CTAS];
#pragma unroll
            for (int u = 0; u < GROUP_CTAS; ++u)
                t[u] = b0 + u < b1 ? __ldcg(src + size_t(b0 + u) * ng4) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            float4 acc = t[0];
#pragma unroll
            for (int u = 1; u < GROUP_CTAS; ++u) {
                acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w;
            }
            reinterpret_cast<float4*>(gpart)[size_t(mygroup) * ng4 + g] = acc;
        }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s.role = atomicAdd(&tick[TICKET_TOP], 1u) == ngroups - 1 ? 2 : 1;
    __syncthreads();
    WX_TC_STAMP(5);
    if (s.role != 2) return;
    __threadfence();
    // level 2: the group sums in group order (all of a thread's loads in flight at once), then the update.  Without a
    // peer exchange every thread updates the four parameters it has just summed -- their p/m/v were requested before the
    // sums -- and nothing goes through shared memory.
    constexpr int MAXG = (148 + GROUP_CTAS - 1) / GROUP_CTAS;         // 13 groups at most on a 148-SM part
    const bool exchange = a.peer.world > 1 && a.tail == wxk::TAIL_ADAM;
    const bool vec_ok = a.tail == wxk::TAIL_ADAM && !exchange &&
                        ((reinterpret_cast<uintptr_t>(a.p_rw) | reinterpret_cast<uintptr_t>(a.m) | reinterpret_cast<uintptr_t>(a.v)) & 15) == 0;
    float* bucket = s.u[1];
    bool apply = true;
    for (int g = tid; g < ng4; g += TC_THREADS) {
        // vec_ok: every thread updates the (up to) four parameters it sums, in this same pass -- the group that straddles P
        // element by element, with the mean loss in the element after the last parameter
        const bool direct = vec_ok && 4 * g < P + 1;
        const bool full = 4 * g + 3 < P;
        float pq[4] = {0.0f, 0.0f, 0.0f, 0.0f}, mq[4] = {0.0f, 0.0f, 0.0f, 0.0f}, vq[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (direct && full) {
            const float4 p4 = reinterpret_cast<const float4*>(a.p_rw)[g], m4 = reinterpret_cast<const float4*>(a.m)[g],
                         v4 = reinterpret_cast<const float4*>(a.v)[g];
            pq[0] = p4.x; pq[1] = p4.y; pq[2] = p4.z; pq[3] = p4.w;
            mq[0] = m4.x; mq[1] = m4.y; mq[2] = m4.z; mq[3] = m4.w;
            vq[0] = v4.x; vq[1] = v4.y; vq[2] = v4.z; vq[3] = v4.w;
        } else if (direct) {
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (4 * g + e < P) {
                    pq[e] = a.p_rw[4 * g + e];
                    mq[e] = a.m[4 * g + e];
                    vq[e] = a.v[4 * g + e];
                }
        }
        const float4* src = reinterpret_cast<const float4*>(gpart) + g;
        float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (ngroups <= MAXG) {
            float4 t[MAXG];
#pragma unroll
            for (int u = 0; u < MAXG; ++u)
                t[u] = unsigned(u) < ngroups ? __ldcg(src + size_t(u) * ng4) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
            for (int u = 0; u < MAXG; ++u) {
                acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w;
            }
        } else {
            for (unsigned u = 0; u < ngroups; ++u) {
                const float4 t = __ldcg(src + size_t(u) * ng4);
                acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
            }
        }
        if (direct) {
            const float gq[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (full || 4 * g + e < P) wxk::adam_apply(pq[e], gq[e], mq[e], vq[e], s.adam);
            if (full) {
                reinterpret_cast<float4*>(a.p_rw)[g] = make_float4(pq[0], pq[1], pq[2], pq[3]);
                reinterpret_cast<float4*>(a.m)[g] = make_float4(mq[0], mq[1], mq[2], mq[3]);
                reinterpret_cast<float4*>(a.v)[g] = make_float4(vq[0], vq[1], vq[2], vq[3]);
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int i = 4 * g + e;
                    if (i < P) {
                        a.p_rw[i] = pq[e];
                        a.m[i] = mq[e];
                        a.v[i] = vq[e];
                    } else if (i == P) {
                        *a.loss_out = gq[e] * a.scale;                // mean loss
                    }
                }
            }
        } else {
            reinterpret_cast<float4*>(bucket)[g] = acc;               // gradients out, or DDP: through the bucket
        }
    }
    if (!vec_ok) {
        __syncthreads();
        if (tid == 0) bucket[P] *= a.scale;                           // mean loss
        __syncthreads();
        if (exchange) {
            // DDP exchange inside the kernel: one-shot all-reduce over NVLink peer memory (csrc/peer.cuh).  If a peer does
            // not show up in time the step is NOT applied on this rank (parameters, moments and step counter stay as they
            // were) and the sticky error code says so -- a visible stall instead of silently diverging replicas.
            apply = wxpeer::allreduce_cta(bucket, P + 1, a.peer, uint64_t(a.state[0]) + 1, tid, TC_THREADS);
            if (!apply && tid == 0) atomicExch(err_out, 20);
        }
        for (int i = tid; i < P + 1; i += TC_THREADS) {
            const float sum = bucket[i];
            if (a.tail == wxk::TAIL_GRADS) {
                a.out[i] = sum;
            } else if (!apply) {
                continue;
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
    }
    WX_TC_STAMP(6);
    if (tid < 32) {
        for (unsigned g = tid; g < ngroups; g += 32) tick[TICKET_GROUP0 + g] = 0u;
        if (tid == 0) {
            tick[TICKET_TOP] = 0u;
            if (a.tail == wxk::TAIL_ADAM && apply) a.state[0] = int64_t(step_word) + 1;   // read at kernel start: a store, no round trip
        }
    }
}

template <bool RAW, int DROP, int FT>
int tc_launch_inst(wx_ctx* ctx, const GridArgs& a, unsigned grid, int smem, const TcGeom& geo, int* err, cudaStream_t st) {
    // the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute
    static bool attr_set[64] = {};
    const int dev = ctx->device;
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        WX_CUDA(cudaFuncSetAttribute(tc_step_kernel<RAW, DROP, FT>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    // programmatic dependent launch: this launch's prologue (barrier init, TMEM allocation) may overlap the tail of the
    // previous kernel in the stream; the kernel itself waits (griddepcontrol.wait) before it reads anything a
    // predecessor may have written
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = size_t(smem);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    WX_CUDA(cudaLaunchKernelEx(&cfg, tc_step_kernel<RAW, DROP, FT>, a, ctx->desc.n_features, geo, err));
    return WX_OK;
}

template <int FT>
int tc_launch(wx_ctx* ctx, GridArgs& a, cudaStream_t st, const char* who) {
    const int f = ctx->desc.n_features;
    WX_CHECK_ARG(FT == 0 || FT == f, "%s: kernel built for F=%d launched with F=%d", who, FT, f);
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
    // landing ring: as many stages (at most 3) as fit beside the fixed operand tiles
    TcGeom geo;
    geo.x_bytes = (TILE * f * 4 + 127) & ~127;
    geo.stage_bytes = geo.x_bytes + 2 * TILE * 8;
    const int fixed = int(offsetof(TcSmem, landing));
    geo.stages = (SMEM_LIMIT - fixed) / geo.stage_bytes;
    if (geo.stages > MAX_STAGES) geo.stages = MAX_STAGES;
    if (geo.stages < 2) {
        wx_set_error("%s: shared memory does not hold two landing stages for n_features=%d", who, f);
        return WX_ENOMEM;
    }
    const int smem = fixed + geo.stages * geo.stage_bytes;
    const int64_t ntiles = (a.batch + TILE - 1) / TILE;
    int64_t grid = ntiles < ctx->num_sms ? ntiles : ctx->num_sms;
    const size_t slab = (size_t(64) * f + 194 + 1 + 3) & ~size_t(3);
    const size_t ngroups = (size_t(grid) + GROUP_CTAS - 1) / GROUP_CTAS;
    if (WX_SCRATCH_PART_OFF + (size_t(grid) + ngroups) * slab * 4 > ctx->scratch_bytes ||
        ngroups > size_t(TICKET_TOP - TICKET_GROUP0)) {
        wx_set_error("%s: scratch too small", who);
        return WX_ENOMEM;
    }
    a.part = reinterpret_cast<float*>(ctx->scratch + WX_SCRATCH_PART_OFF);
    a.ticket = reinterpret_cast<unsigned int*>(ctx->scratch) + 1;
    a.desc = ctx->desc;
#ifdef WX_TC_TIMERS
    a.dbg = ctx->debug_times;      // developer build: per-CTA globaltimer stamps [grid][8]
#else
    a.dbg = nullptr;
#endif
    if (a.tail == wxk::TAIL_ADAM && ctx->peer.world > 1) {
        a.peer = ctx->peer;
        a.scale = a.scale / float(ctx->peer.world);          // mean over W*B rows (what DDP's averaged gradients give)
    }
    int* err = reinterpret_cast<int*>(ctx->scratch) + 8;       // word 8 of the ticket block: sticky protocol-error code
    const bool rawl = a.loader != wxk::LOAD_ROWMAJOR;
    int rc;
    const unsigned g = unsigned(grid);
    switch (a.dropout_mode) {
        case WX_DROPOUT_OFF:
            rc = rawl ? tc_launch_inst<true, DROP_OFF, FT>(ctx, a, g, smem, geo, err, st)
                      : tc_launch_inst<false, DROP_OFF, FT>(ctx, a, g, smem, geo, err, st);
            break;
        case WX_DROPOUT_MASK:
            rc = rawl ? tc_launch_inst<true, DROP_MASK, FT>(ctx, a, g, smem, geo, err, st)
                      : tc_launch_inst<false, DROP_MASK, FT>(ctx, a, g, smem, geo, err, st);
            break;
        default:
            rc = rawl ? tc_launch_inst<true, DROP_PHILOX, FT>(ctx, a, g, smem, geo, err, st)
                      : tc_launch_inst<false, DROP_PHILOX, FT>(ctx, a, g, smem, geo, err, st);
            break;
    }
    if (rc != WX_OK) return rc;
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}


}  // namespace wxtcstep
