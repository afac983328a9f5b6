// bf16 tcgen05 GEMM for the wide tabular MLP (BASELINE.json configs[3]): D[M x N] = A * B with fp32
// accumulation in TMEM.  Not in the reference (its model is 5->64->2, /root/reference/jobs/train_lightning_ddp.py:57-62);
// this is the "wide-hidden config where the layers are real dense GEMMs" of the north star.
//
// One persistent CTA per SM, 352 threads, warp-specialised:
//   warp 0      TMA producer for A, warp 1 TMA producer for B: tensor-map loads (128B swizzle) of a 128 x 64 A block and
//               a 256 x 64 B block per stage, ONE bulk-tensor op per operand and stage (a 3-D map walks the 64-column
//               atoms of an MN-major operand)
//   warp 2      MMA issuer: 4 x tcgen05.mma.kind::f16 (M128 N256 K16) per stage, tcgen05.commit frees the stage
//   warps 3..10 epilogue: tcgen05.ld of the finished accumulator (TMEM is double-buffered: 2 x 256 columns, so the
//               epilogue of tile i overlaps the main loop of tile i+1), fused pointwise work, global stores; two warps
//               per TMEM lane quarter, each taking half of the 256 columns
// Operand layouts in global memory are all row-major; an operand whose contraction index is the row index (the W of the
// dX product, both operands of the dW product) is read "MN-major": TMA boxes of 64 x 64 land as the canonical MN-major
// 128B-swizzle atoms and the UMMA descriptor says so, so no transposed copies of activations or weights exist anywhere.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace wxwide {

constexpr int BM = 128, BN = 256, BK = 64, STAGES = 4;
constexpr int A_STAGE_BYTES = BM * BK * 2;   // 16 KiB
constexpr int B_STAGE_BYTES = BN * BK * 2;   // 32 KiB
constexpr int GEMM_THREADS = 352;
constexpr int EPI_WARPS = 8;
constexpr size_t GEMM_SMEM = size_t(STAGES) * (A_STAGE_BYTES + B_STAGE_BYTES) + 1024;   // + alignment slack

enum Epilogue : int {
    EPI_FWD = 0,    // out bf16 = dropout(relu(acc + bias[col]))           (train_lightning_ddp.py:58-60 pattern per layer)
    EPI_DX = 1,     // out bf16 = acc * (gate[row][col] > 0 ? 1/keep : 0)   (backward through dropout + ReLU)
    EPI_DW = 2,     // out fp32 = acc                                       (weight gradient, written once per element)
    EPI_PLAIN = 3,  // out fp32 = acc (+ bias if given)                     (self-test)
};

struct GemmArgs {
    int64_t M, N, K;          // output M x N, contraction length K (multiple of 64)
    int split_k;              // >= 1: the contraction is cut into this many ranges, one output slab each (EPI_DW / PLAIN)
    int64_t split_stride;     // elements between the output slabs of consecutive splits
    void* out;
    int64_t ldo;              // output row pitch in elements
    const float* bias;        // [N] or NULL
    const __nv_bfloat16* gate;
    int64_t ldg;
    float inv_keep;           // 1 / (1 - p); 1 when dropout is off
    uint32_t keep_thresh16;   // keep <=> 16 random bits < this; 65536 when dropout is off
    uint32_t seed_lo, seed_hi, step, layer;
    int* err;
    int nb;                   // two-SM kernel: 0 = pick the tile width, 1 = 256 x 256, 2 = 256 x 512
    // Tail split (two-SM kernel, split_k == 1, fp32 output): output tiles [0, tail_first) are whole work units, every later
    // tile is cut into tail_split ranges of K -- range 0 goes to `out`, range s >= 1 to out + tail_off + (s-1) * tail_stride
    // -- so that the last, partly filled round of tiles is spread over all SM pairs.  tail_split <= 1: off.
    int tail_split;
    int64_t tail_first, tail_off, tail_stride;
    float* colsum_part;       // EPI_DX only, may be NULL: [ceil(M/32)][N] column sums of each 32-row block of the bf16 OUTPUT
                              // (the bias gradient of the layer below falls out of the dZ product that writes its dZ)
};

// ---- Philox4x32-10, 16 bits per element: the wide path's dropout contract (oracle/wide_ref.py: keep_mask) -------------
__device__ __forceinline__ void philox4(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                        uint32_t (&r)[4]) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        const uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        c0 = h1 ^ c1 ^ k0;
        c1 = l1;
        c2 = h0 ^ c3 ^ k1;
        c3 = l0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    r[0] = c0; r[1] = c1; r[2] = c2; r[3] = c3;
}

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tm, int32_t c0, int32_t c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(wxtc::smem_u32(smem_dst)), "l"(tm), "r"(c0), "r"(c1), "r"(wxtc::smem_u32(bar))
        : "memory");
}

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* tm, int32_t c0, int32_t c1, int32_t c2,
                                            uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(wxtc::smem_u32(smem_dst)), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(wxtc::smem_u32(bar))
        : "memory");
}

__host__ __device__ constexpr uint32_t make_idesc_bf16(int m, int n, bool a_mn_major, bool b_mn_major) {
    return (1u << 4)                       // c_format = F32
           | (1u << 7)                     // a_format = BF16
           | (1u << 10)                    // b_format = BF16
           | (uint32_t(a_mn_major) << 15)  // a_major
           | (uint32_t(b_mn_major) << 16)  // b_major
           | (uint32_t(n >> 3) << 17)      // n_dim
           | (uint32_t(m >> 4) << 24);     // m_dim
}

__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    const uint32_t acc = accumulate ? 1u : 0u;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

// Sums of 16 per-lane values over the 32 lanes of a warp with one transposing butterfly (8 + 4 + 2 + 1 + 1 shuffles):
// afterwards lanes 2k and 2k+1 both hold the total of value `col` (returned per lane: col = 8 b4 + 4 b3 + 2 b2 + b1 of the
// lane index).  Fixed order, so the result is reproducible.
__device__ __forceinline__ float warp_transpose_sum16(const float (&r)[16], int lane, int& col) {
    float u8[8], u4[4], u2[2];
    {
        const bool hi = lane & 16;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float send = hi ? r[i] : r[i + 8];
            const float keep = hi ? r[i + 8] : r[i];
            u8[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool hi = lane & 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float send = hi ? u8[i] : u8[i + 4];
            const float keep = hi ? u8[i + 4] : u8[i];
            u4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool hi = lane & 4;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const float send = hi ? u4[i] : u4[i + 2];
            const float keep = hi ? u4[i + 2] : u4[i];
            u2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
    }
    const bool hi2 = lane & 2;
    float t = (hi2 ? u2[1] : u2[0]) + __shfl_xor_sync(0xffffffffu, hi2 ? u2[0] : u2[1], 2);
    t += __shfl_xor_sync(0xffffffffu, t, 1);
    col = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
    return t;
}

// Fused pointwise work on one 16-column chunk of an accumulator row (shared by the 1-CTA and 2-CTA kernels).  gw: the 16
// bf16 gate values of this row and chunk (EPI_DX only).
template <int EPI>
__device__ __forceinline__ void epilogue_math(float (&v)[16], const GemmArgs& g, int64_t row, bool row_ok, int64_t col0,
                                              const uint32_t (&gw)[8]) {
    if (EPI == EPI_FWD) {
        uint32_t keep = 0xFFFFu;
        if (g.keep_thresh16 < 65536u) {
            keep = 0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                uint32_t r[4];
                philox4(uint32_t(row), uint32_t((col0 >> 3) + h), g.layer, g.step, g.seed_lo, g.seed_hi, r);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const uint32_t u = (r[e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
                    keep |= uint32_t(u < g.keep_thresh16) << (h * 8 + e);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            float z = v[i] + (g.bias ? __ldg(g.bias + col0 + i) : 0.0f);
            z = fmaxf(z, 0.0f);
            v[i] = ((keep >> i) & 1u) ? z * g.inv_keep : 0.0f;
        }
    } else if (EPI == EPI_DX) {
        if (row_ok) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const uint32_t bits = (gw[i >> 1] >> (16 * (i & 1))) & 0xFFFFu;
                const bool on = (bits & 0x7FFFu) != 0 && !(bits & 0x8000u);   // gate > 0
                v[i] = on ? v[i] * g.inv_keep : 0.0f;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = 0.0f;
        }
        if (g.colsum_part != nullptr) {
            // column sums over this warp's 32 rows of the values AS STORED (bf16-rounded): one transposing butterfly
            // (8 + 4 + 2 + 1 + 1 shuffles), after which lanes 2k and 2k+1 hold the total of column k
            float r[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = __bfloat162float(__float2bfloat16_rn(v[i]));
            int col;
            const float t = warp_transpose_sum16(r, threadIdx.x & 31, col);
            const int lane = threadIdx.x & 31;
            if ((lane & 1) == 0) g.colsum_part[(row >> 5) * g.N + col0 + col] = t;
        }
    } else if (EPI == EPI_PLAIN) {
        if (g.bias) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] += __ldg(g.bias + col0 + i);
        }
    }
}

// epilogue_math + the store, straight from registers: every lane writes 32 or 64 bytes of its own row (the 1-CTA kernel)
template <int EPI>
__device__ __forceinline__ void epilogue_chunk(float (&v)[16], const GemmArgs& g, int64_t row, bool row_ok, int64_t col0,
                                               int64_t out_off) {
    uint32_t gw[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (EPI == EPI_DX && row_ok) {
        const uint4* gp = reinterpret_cast<const uint4*>(g.gate + row * g.ldg + col0);
        const uint4 g0 = __ldg(gp), g1 = __ldg(gp + 1);
        gw[0] = g0.x; gw[1] = g0.y; gw[2] = g0.z; gw[3] = g0.w;
        gw[4] = g1.x; gw[5] = g1.y; gw[6] = g1.z; gw[7] = g1.w;
    }
    epilogue_math<EPI>(v, g, row, row_ok, col0, gw);
    if (row_ok) {
        if (EPI == EPI_FWD || EPI == EPI_DX) {
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(g.out) + row * g.ldo + col0;
            uint4 o0, o1;
            o0.x = pack_bf16(v[0], v[1]);   o0.y = pack_bf16(v[2], v[3]);
            o0.z = pack_bf16(v[4], v[5]);   o0.w = pack_bf16(v[6], v[7]);
            o1.x = pack_bf16(v[8], v[9]);   o1.y = pack_bf16(v[10], v[11]);
            o1.z = pack_bf16(v[12], v[13]); o1.w = pack_bf16(v[14], v[15]);
            reinterpret_cast<uint4*>(o)[0] = o0;
            reinterpret_cast<uint4*>(o)[1] = o1;
        } else {
            float* o = reinterpret_cast<float*>(g.out) + out_off + row * g.ldo + col0;
#pragma unroll
            for (int i = 0; i < 4; ++i)
                reinterpret_cast<float4*>(o)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
    }
}

// A_K: A is [M x K] row-major (contraction contiguous); else A is [K x M] row-major.  Same for B with N.
template <bool A_K, bool B_K, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
wide_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmArgs g) {
    extern __shared__ unsigned char wide_smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(wide_smem_raw) + 1023) & ~uintptr_t(1023));
    unsigned char* sA = smem;
    unsigned char* sB = smem + size_t(STAGES) * A_STAGE_BYTES;
    __shared__ uint64_t full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tiles_m = (g.M + BM - 1) / BM, tiles_n = (g.N + BN - 1) / BN;
    const int64_t mn_tiles = tiles_m * tiles_n;
    const int64_t n_tiles = mn_tiles * g.split_k;
    const int kb_total = int(g.K / BK);
    const int kb_per = (kb_total + g.split_k - 1) / g.split_k;   // k-blocks per split (the last one may be shorter)

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            wxtc::mbar_init(&full[s], 2);      // one arrive.expect_tx from each producer
            wxtc::mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            wxtc::mbar_init(&tmem_full[a], 1);
            wxtc::mbar_init(&tmem_empty[a], EPI_WARPS);
        }
        wxtc::fence_mbar_init();
    }
    if (warp == 2) wxtc::tmem_alloc(&tmem_base_s, 512);
    wxtc::tc_fence_before();
    __syncthreads();
    wxtc::tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    bool ok = true;

    if (warp < 2) {
        // ================= TMA producers: warp 0 loads A, warp 1 loads B =================
        if (lane == 0) {
            const bool is_a = warp == 0;
            uint32_t it = 0;   // running stage counter across tiles
            for (int64_t tile = blockIdx.x; tile < n_tiles && ok; tile += gridDim.x) {
                const int64_t mn = tile % mn_tiles;
                const int split = int(tile / mn_tiles);
                const int32_t m0 = int32_t(mn / tiles_n) * BM, n0 = int32_t(mn % tiles_n) * BN;
                const int kb0 = split * kb_per, kb1 = (kb0 + kb_per < kb_total) ? kb0 + kb_per : kb_total;
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int s = it % STAGES;
                    if (it >= STAGES && !wxtc::mbar_wait_spin(&empty[s], ((it / STAGES) - 1) & 1)) { ok = false; break; }
                    if (is_a) {
                        wxtc::mbar_arrive_expect_tx(&full[s], A_STAGE_BYTES);
                        unsigned char* a = sA + size_t(s) * A_STAGE_BYTES;
                        if (A_K) tma_load_2d(a, &tmA, kb * BK, m0, &full[s]);
                        else tma_load_3d(a, &tmA, 0, kb * BK, m0 / 64, &full[s]);
                    } else {
                        wxtc::mbar_arrive_expect_tx(&full[s], B_STAGE_BYTES);
                        unsigned char* b = sB + size_t(s) * B_STAGE_BYTES;
                        if (B_K) tma_load_2d(b, &tmB, kb * BK, n0, &full[s]);
                        else tma_load_3d(b, &tmB, 0, kb * BK, n0 / 64, &full[s]);
                    }
                }
            }
            if (!ok) atomicExch(g.err, 201);
        }
    } else if (warp == 2) {
        // ================= MMA issuer =================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(BM, BN, !A_K, !B_K);
            uint32_t it = 0, local = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles && ok; tile += gridDim.x, ++local) {
                const uint32_t acc = local & 1;
                if (local >= 2 && !wxtc::mbar_wait_spin(&tmem_empty[acc], ((local >> 1) - 1) & 1)) { ok = false; break; }
                wxtc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                const int split = int(tile / mn_tiles);
                const int kb0 = split * kb_per, kb1 = (kb0 + kb_per < kb_total) ? kb0 + kb_per : kb_total;
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int s = it % STAGES;
                    if (!wxtc::mbar_wait_spin(&full[s], (it / STAGES) & 1)) { ok = false; break; }
                    wxtc::tc_fence_after();
                    const uint32_t a_addr = wxtc::smem_u32(sA + size_t(s) * A_STAGE_BYTES);
                    const uint32_t b_addr = wxtc::smem_u32(sB + size_t(s) * B_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        // K-major: 16 bf16 = 32 bytes further along the 128-byte line; 8-row groups 1024 B apart.
                        // MN-major: 16 k-lines = 2048 bytes further; 64-element M|N atoms 8192 B apart, 8-k groups 1024 B.
                        const uint64_t ad = A_K ? wxtc::make_smem_desc(a_addr + k * 32, 16, 1024)
                                                : wxtc::make_smem_desc(a_addr + k * 2048, 8192, 1024);
                        const uint64_t bd = B_K ? wxtc::make_smem_desc(b_addr + k * 32, 16, 1024)
                                                : wxtc::make_smem_desc(b_addr + k * 2048, 8192, 1024);
                        umma_bf16(d_tmem, ad, bd, idesc, ((kb - kb0) | k) != 0);
                    }
                    wxtc::umma_commit(&empty[s]);          // stage reusable once these MMAs have read it
                }
                if (!ok) break;
                wxtc::umma_commit(&tmem_full[acc]);         // accumulator complete
            }
            if (!ok) atomicExch(g.err, 202);
        }
    } else {
        // ================= epilogue warps (TMEM lane quarter = warp % 4; column half = (warp - 3) / 4) =================
        const int q = warp & 3;
        const int half = (warp - 3) >> 2;
        uint32_t local = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++local) {
            const uint32_t acc = local & 1;
            const int64_t mn = tile % mn_tiles;
            const int64_t m0 = (mn / tiles_n) * BM, n0 = (mn % tiles_n) * BN;
            const int64_t out_off = (tile / mn_tiles) * g.split_stride;
            if (!wxtc::mbar_wait(&tmem_full[acc], (local >> 1) & 1)) {
                if (lane == 0) atomicExch(g.err, 203);
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
                const int64_t col0 = n0 + c * 16;
                if (col0 >= g.N) break;                      // warp-uniform
                epilogue_chunk<EPI>(v, g, row, row_ok, col0, out_off);
                __syncwarp();
            }
            wxtc::tc_fence_before();
            __syncwarp();
            if (lane == 0) wxtc::mbar_arrive(&tmem_empty[acc]);
        }
    }
    wxtc::tc_fence_before();
    __syncthreads();
    if (warp == 2) wxtc::tmem_dealloc(tmem_base, 512);
}

}  // namespace wxwide
