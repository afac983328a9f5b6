// Wide tabular MLP (F -> H -> ... -> H -> 2, H a multiple of 256; BASELINE.json configs[3]: 4 layers, hidden 4096) on the
// bf16 tcgen05 GEMM of wide_gemm.cuh.  Same layer pattern as the reference's net (Linear, ReLU, Dropout(p) per hidden
// layer, Linear head, CE loss, Adam -- /root/reference/jobs/train_lightning_ddp.py:57-62,:69,:88), widened; the reference
// has no such model, so the oracle for this path is oracle/wide_ref.py (torch fp32 with bf16 rounding at the same points).
//
// Data flow of one step (all device pointers caller-owned; `ws` is one caller-allocated workspace):
//   x fp32 [B,F] --cast/pad--> X0 bf16 [B,64]
//   H_l = drop(relu(H_{l-1} W_l^T + b_l))            GEMM, EPI_FWD      l = 1..L   (bf16 activations, fp32 accumulate)
//   logits = H_L W_o^T + b_o ; loss ; dlogits ; dZ_L = (dlogits W_o) * gate(H_L)
//                                                      wide_head_rows_kernel (hidden <= 4096; else wide_head_kernel)
//   dW_o partials                                      wide_head_bwd_kernel
//   for l = L..1:  dW_l = dZ_l^T H_{l-1}              GEMM, EPI_DW (both operands read MN-major, fp32 out)
//                  db_l = colsum(dZ_l)                 top layer: wide_colsum_kernel + wide_reduce_kernel; below: 32-row
//                                                      partials from the EPI_DX epilogue + wide_reduce_kernel
//                  dZ_{l-1} = (dZ_l W_l) * gate(H_{l-1})   GEMM, EPI_DX (W read MN-major)
//   Adam on the fp32 master parameters, re-emitting the bf16 copies of the hidden weights in the same pass.
#include "wide_gemm2.cuh"

#include <stdlib.h>

#include <math.h>

namespace {

using namespace wxwide;

struct AdamS {
    float w_lerp, beta2, one_m_beta2, neg_step, bc2_sqrt, eps;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamS& a) {
    // torch.optim.Adam single-tensor op order (same as wxk::adam_apply)
    m = fmaf(a.w_lerp, g - m, m);
    v = v * a.beta2;
    v = v + (a.one_m_beta2 * g) * g;
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
    p = p + (a.neg_step * m) / denom;
}

constexpr int KPAD = 64;          // the first layer's contraction length, zero padded (F <= 64)
constexpr int COLSUM_SLABS = 128;
constexpr int HEAD_SLABS = 128;
constexpr int MAX_SPLIT_K = 8;
constexpr int MAX_TAIL_SPLIT = 4;

struct Layout {
    int64_t P;                    // number of fp32 parameters
    int64_t off_w[8], off_b[8];   // parameter offsets of hidden layer l = 1..L (index l), head at index L+1
    size_t wb[8];                 // workspace byte offsets: bf16 weights of hidden layer l
    size_t x0, h[8], dz[2], dlog, part, part2, dw1tmp, dwtmp[8], cpart[8], total;
};

size_t align_up(size_t v) { return (v + 1023) & ~size_t(1023); }

int make_layout(const wx_wide_desc* d, int64_t batch, Layout* L) {
    const int64_t F = d->n_features, H = d->hidden, C = d->classes;
    const int nl = d->n_hidden;
    int64_t off = 0;
    for (int l = 1; l <= nl; ++l) {
        const int64_t in = (l == 1) ? F : H;
        L->off_w[l] = off;
        off += H * in;
        L->off_b[l] = off;
        off += H;
    }
    L->off_w[nl + 1] = off;
    off += C * H;
    L->off_b[nl + 1] = off;
    off += C;
    L->P = off;
    size_t w = 0;
    for (int l = 1; l <= nl; ++l) {
        L->wb[l] = w;
        w = align_up(w + size_t(H) * ((l == 1) ? KPAD : H) * 2);
    }
    L->x0 = w;
    w = align_up(w + size_t(batch) * KPAD * 2);
    for (int l = 1; l <= nl; ++l) {
        L->h[l] = w;
        w = align_up(w + size_t(batch) * H * 2);
    }
    for (int i = 0; i < 2; ++i) {
        L->dz[i] = w;
        w = align_up(w + size_t(batch) * H * 2);
    }
    L->dlog = w;
    w = align_up(w + size_t(batch) * 2 * 4);
    L->part = w;   // partial sums: max(colsum: SLABS x H, head: SLABS x (2H), head rows: grid x 4)
    w = align_up(w + size_t(HEAD_SLABS) * 2 * H * 4 + size_t(4096) * 4 * 4);
    L->part2 = w;   // head dW_o partials (their own buffer: that piece may run beside the column sums)
    w = align_up(w + size_t(HEAD_SLABS) * 2 * H * 4);
    L->dw1tmp = w;
    w = align_up(w + size_t(MAX_SPLIT_K) * H * KPAD * 4);
    for (int l = 2; l <= nl; ++l) {   // K-range slabs 1..3 of dW_l's tail tiles (wave quantisation: see plan_dw)
        L->dwtmp[l] = w;
        w = align_up(w + size_t(MAX_TAIL_SPLIT - 1) * H * H * 4);
    }
    for (int l = 1; l < nl; ++l) {    // 32-row column sums of dZ_l written by the dZ product of layer l+1 (bias gradient of l)
        L->cpart[l] = w;
        w = align_up(w + size_t((batch + 31) / 32) * H * 4);
    }
    L->total = w;
    return 0;
}

int check_desc(const wx_wide_desc* d, const char* who) {
    WX_CHECK_ARG(d != nullptr, "%s: desc is NULL", who);
    WX_CHECK_ARG(d->n_features >= 1 && d->n_features <= KPAD, "%s: n_features=%d outside 1..64", who, d->n_features);
    WX_CHECK_ARG(d->hidden >= 256 && d->hidden % 256 == 0 && d->hidden <= 16384, "%s: hidden=%d must be a multiple of 256 in 256..16384",
                 who, d->hidden);
    WX_CHECK_ARG(d->n_hidden >= 1 && d->n_hidden <= 6, "%s: n_hidden=%d outside 1..6", who, d->n_hidden);
    WX_CHECK_ARG(d->classes == 2, "%s: classes=%d, only 2 is supported (train_lightning_ddp.py:61)", who, d->classes);
    WX_CHECK_ARG(d->dropout_p >= 0.0f && d->dropout_p < 1.0f, "%s: dropout_p=%f outside [0,1)", who, d->dropout_p);
    return WX_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// small kernels
// ------------------------------------------------------------------------------------------------------------------
// fp32 [rows, in] -> bf16 [rows, pad] (zero padded); used for x and for the first layer's weights
__global__ void __launch_bounds__(256) cast_pad_kernel(const float* __restrict__ src, int64_t rows, int in, int pad,
                                                      __nv_bfloat16* __restrict__ dst) {
    const int64_t total = rows * pad;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += int64_t(gridDim.x) * blockDim.x) {
        const int64_t r = i / pad;
        const int c = int(i - r * pad);
        dst[i] = __float2bfloat16_rn(c < in ? src[r * in + c] : 0.0f);
    }
}

__global__ void __launch_bounds__(256) cast_kernel(const float* __restrict__ src, int64_t n, __nv_bfloat16* __restrict__ dst) {
    const int64_t n4 = n >> 2;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += int64_t(gridDim.x) * blockDim.x) {
        const float4 v = reinterpret_cast<const float4*>(src)[i];
        uint2 o;
        o.x = pack_bf16(v.x, v.y);
        o.y = pack_bf16(v.z, v.w);
        reinterpret_cast<uint2*>(dst)[i] = o;
    }
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i] = __uint_as_float(w[i] << 16);
        f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
}

// head forward + loss: one warp per row.  logits = h . W_o^T + b_o (fp32), CE, dlogits = (softmax - onehot) * scale.
// Per-CTA partial (loss*scale, sum dl0, sum dl1) -> part[blockIdx][4]; reduced in order by wide_reduce_kernel.
__global__ void __launch_bounds__(256) wide_head_kernel(const __nv_bfloat16* __restrict__ h, int64_t batch, int H,
                                                       const float* __restrict__ w_o, const float* __restrict__ b_o,
                                                       const int64_t* __restrict__ y, float scale,
                                                       float* __restrict__ dlog, float* __restrict__ part,
                                                       float inv_keep, __nv_bfloat16* __restrict__ dz) {
    // dz != NULL: the same warp also writes dZ_L[r][c] = (dl0 W_o[0][c] + dl1 W_o[1][c]) * gate(H_L[r][c]) for its row,
    // re-reading the row it has just pulled through L1/L2 -- H_L crosses DRAM once for logits AND dZ.
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float acc_loss = 0.0f, acc0 = 0.0f, acc1 = 0.0f, acc_ok = 0.0f;
    for (int64_t row = int64_t(blockIdx.x) * 8 + warp; row < batch; row += int64_t(gridDim.x) * 8) {
        const uint4* hp = reinterpret_cast<const uint4*>(h + row * H);
        float s0 = 0.0f, s1 = 0.0f;
        for (int c8 = lane; c8 < H / 8; c8 += 32) {
            float f[8];
            unpack8(__ldg(hp + c8), f);
            const float4* w0 = reinterpret_cast<const float4*>(w_o + c8 * 8);
            const float4* w1 = reinterpret_cast<const float4*>(w_o + H + c8 * 8);
            const float4 a0 = __ldg(w0), a1 = __ldg(w0 + 1), c0 = __ldg(w1), c1 = __ldg(w1 + 1);
            s0 = fmaf(f[0], a0.x, s0); s0 = fmaf(f[1], a0.y, s0); s0 = fmaf(f[2], a0.z, s0); s0 = fmaf(f[3], a0.w, s0);
            s0 = fmaf(f[4], a1.x, s0); s0 = fmaf(f[5], a1.y, s0); s0 = fmaf(f[6], a1.z, s0); s0 = fmaf(f[7], a1.w, s0);
            s1 = fmaf(f[0], c0.x, s1); s1 = fmaf(f[1], c0.y, s1); s1 = fmaf(f[2], c0.z, s1); s1 = fmaf(f[3], c0.w, s1);
            s1 = fmaf(f[4], c1.x, s1); s1 = fmaf(f[5], c1.y, s1); s1 = fmaf(f[6], c1.z, s1); s1 = fmaf(f[7], c1.w, s1);
        }
        s0 = warp_sum(s0) + b_o[0];
        s1 = warp_sum(s1) + b_o[1];
        // every lane evaluates the (identical) softmax so that all of them hold the dlogits for the dZ pass
        const float mx = fmaxf(s0, s1);
        const float e0 = expf(s0 - mx), e1 = expf(s1 - mx);
        const float lse = mx + logf(e0 + e1);
        const int64_t lab = y[row];
        const float inv = 1.0f / (e0 + e1);
        const float d0 = (e0 * inv - (lab == 0 ? 1.0f : 0.0f)) * scale;
        const float d1 = (e1 * inv - (lab == 1 ? 1.0f : 0.0f)) * scale;
        if (dz != nullptr) {
            uint4* zp = reinterpret_cast<uint4*>(dz + row * H);
            for (int c8 = lane; c8 < H / 8; c8 += 32) {
                float f[8], o[8];
                unpack8(__ldg(hp + c8), f);
                const float4* w0 = reinterpret_cast<const float4*>(w_o + c8 * 8);
                const float4* w1 = reinterpret_cast<const float4*>(w_o + H + c8 * 8);
                const float4 a0 = __ldg(w0), a1 = __ldg(w0 + 1), c0 = __ldg(w1), c1 = __ldg(w1 + 1);
                const float u0[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                const float u1[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float dh = fmaf(d0, u0[i], d1 * u1[i]);
                    o[i] = f[i] > 0.0f ? dh * inv_keep : 0.0f;
                }
                uint4 ov;
                ov.x = pack_bf16(o[0], o[1]);
                ov.y = pack_bf16(o[2], o[3]);
                ov.z = pack_bf16(o[4], o[5]);
                ov.w = pack_bf16(o[6], o[7]);
                zp[c8] = ov;
            }
        }
        if (lane == 0) {
            if (dlog != nullptr) {
                dlog[row * 2] = d0;
                dlog[row * 2 + 1] = d1;
            }
            acc_loss += (lse - (lab == 0 ? s0 : s1)) * scale;
            acc_ok += (int64_t(s1 > s0) == lab) ? 1.0f : 0.0f;          // argmax, first index on ties (torch.argmax)
            acc0 += d0;
            acc1 += d1;
        }
    }
    __shared__ float sh[8][4];
    if (lane == 0) {
        sh[warp][0] = acc_loss;
        sh[warp][1] = acc0;
        sh[warp][2] = acc1;
        sh[warp][3] = acc_ok;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        float t = 0.0f;
        for (int w = 0; w < 8; ++w) t += sh[w][threadIdx.x];
        part[blockIdx.x * 4 + threadIdx.x] = t;
    }
}

// The same head pass for hidden <= 4096, eight rows at a time per CTA: warp w owns the 8-column units k*256 + w*32 + lane of
// every row, so the two head weight rows sit in registers for the whole kernel (the warp-per-row kernel above re-reads
// them from L1 for every row: four weight loads per activation load) and the eight rows' activations stay in registers
// between the logit pass and the dZ pass.  Partial logits: per warp one transposing butterfly over the 16 (row, class)
// sums, then warp r adds the eight warps' parts of row r in warp order.  Same outputs as wide_head_kernel.
template <int K8>
__global__ void __launch_bounds__(256) wide_head_rows_kernel(const __nv_bfloat16* __restrict__ h, int64_t batch, int H,
                                                            const float* __restrict__ w_o, const float* __restrict__ b_o,
                                                            const int64_t* __restrict__ y, float scale,
                                                            float* __restrict__ dlog, float* __restrict__ part,
                                                            float inv_keep, __nv_bfloat16* __restrict__ dz) {
    constexpr int R = 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __shared__ float red[8][2 * R];
    __shared__ float dl[R][2];
    float w0[K8][8], w1[K8][8];
    bool on[K8];
#pragma unroll
    for (int k = 0; k < K8; ++k) {
        const int u = k * 256 + warp * 32 + lane;
        on[k] = u * 8 < H;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            w0[k][i] = on[k] ? __ldg(w_o + u * 8 + i) : 0.0f;
            w1[k][i] = on[k] ? __ldg(w_o + H + u * 8 + i) : 0.0f;
        }
    }
    const float bo0 = b_o[0], bo1 = b_o[1];
    float acc_loss = 0.0f, acc0 = 0.0f, acc1 = 0.0f, acc_ok = 0.0f;      // lane 0 of warp r: row r of every block
    for (int64_t r0 = int64_t(blockIdx.x) * R; r0 < batch; r0 += int64_t(gridDim.x) * R) {
        uint4 hv[R][K8];
#pragma unroll
        for (int r = 0; r < R; ++r) {
#pragma unroll
            for (int k = 0; k < K8; ++k) {
                hv[r][k] = make_uint4(0, 0, 0, 0);
                if (on[k] && r0 + r < batch)
                    hv[r][k] = __ldg(reinterpret_cast<const uint4*>(h + (r0 + r) * H) + k * 256 + warp * 32 + lane);
            }
        }
        float ps[2 * R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
            for (int k = 0; k < K8; ++k) {
                float f[8];
                unpack8(hv[r][k], f);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    s0 = fmaf(f[i], w0[k][i], s0);
                    s1 = fmaf(f[i], w1[k][i], s1);
                }
            }
            ps[2 * r] = s0;
            ps[2 * r + 1] = s1;
        }
        int col;
        const float t = warp_transpose_sum16(ps, lane, col);
        if ((lane & 1) == 0) red[warp][col] = t;
        __syncthreads();
        {
            const int r = warp;                                          // R == number of warps
            float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
            for (int w = 0; w < 8; ++w) {
                s0 += red[w][2 * r];
                s1 += red[w][2 * r + 1];
            }
            s0 += bo0;
            s1 += bo1;
            if (lane == 0 && r0 + r < batch) {
                const float mx = fmaxf(s0, s1);
                const float e0 = expf(s0 - mx), e1 = expf(s1 - mx);
                const float lse = mx + logf(e0 + e1);
                const int64_t lab = y[r0 + r];
                const float inv = 1.0f / (e0 + e1);
                const float d0 = (e0 * inv - (lab == 0 ? 1.0f : 0.0f)) * scale;
                const float d1 = (e1 * inv - (lab == 1 ? 1.0f : 0.0f)) * scale;
                dl[r][0] = d0;
                dl[r][1] = d1;
                if (dlog != nullptr) {
                    dlog[(r0 + r) * 2] = d0;
                    dlog[(r0 + r) * 2 + 1] = d1;
                }
                acc_loss += (lse - (lab == 0 ? s0 : s1)) * scale;
                acc_ok += (int64_t(s1 > s0) == lab) ? 1.0f : 0.0f;      // argmax, first index on ties (torch.argmax)
                acc0 += d0;
                acc1 += d1;
            }
        }
        __syncthreads();
        if (dz != nullptr) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (r0 + r >= batch) break;
                const float d0 = dl[r][0], d1 = dl[r][1];
#pragma unroll
                for (int k = 0; k < K8; ++k) {
                    if (!on[k]) continue;
                    float f[8], o[8];
                    unpack8(hv[r][k], f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float dh = fmaf(d0, w0[k][i], d1 * w1[k][i]);
                        o[i] = f[i] > 0.0f ? dh * inv_keep : 0.0f;
                    }
                    uint4 ov;
                    ov.x = pack_bf16(o[0], o[1]);
                    ov.y = pack_bf16(o[2], o[3]);
                    ov.z = pack_bf16(o[4], o[5]);
                    ov.w = pack_bf16(o[6], o[7]);
                    reinterpret_cast<uint4*>(dz + (r0 + r) * H)[k * 256 + warp * 32 + lane] = ov;
                }
            }
        }
    }
    __shared__ float sh[8][4];
    if (lane == 0) {
        sh[warp][0] = acc_loss;
        sh[warp][1] = acc0;
        sh[warp][2] = acc1;
        sh[warp][3] = acc_ok;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        float t = 0.0f;
        for (int w = 0; w < 8; ++w) t += sh[w][threadIdx.x];
        part[blockIdx.x * 4 + threadIdx.x] = t;
    }
}

// launches the head pass; returns the number of per-CTA partials written to `part`
int launch_head(wx_ctx* ctx, const __nv_bfloat16* h, int64_t batch, int H, const float* w_o, const float* b_o, const int64_t* y,
                float scale, float* dlog, float* part, float inv_keep, __nv_bfloat16* dz, cudaStream_t st) {
    if (H <= 4096) {
        int64_t blocks = (batch + 7) / 8;
        const int64_t cap = int64_t(ctx->num_sms) * 4;
        const int grid = int(blocks < cap ? blocks : cap);
        if (H <= 2048) wide_head_rows_kernel<1><<<grid, 256, 0, st>>>(h, batch, H, w_o, b_o, y, scale, dlog, part, inv_keep, dz);
        else wide_head_rows_kernel<2><<<grid, 256, 0, st>>>(h, batch, H, w_o, b_o, y, scale, dlog, part, inv_keep, dz);
        return grid;
    }
    int hgrid = int((batch + 7) / 8);
    if (hgrid > 4096) hgrid = 4096;
    wide_head_kernel<<<hgrid, 256, 0, st>>>(h, batch, H, w_o, b_o, y, scale, dlog, part, inv_keep, dz);
    return hgrid;
}

// head backward: dZ_L[r][c] = (dl0[r] W_o[0][c] + dl1[r] W_o[1][c]) * gate(H_L[r][c]) as bf16, and per-slab partial sums
// of dW_o[k][c] = sum_r dl_k[r] H_L[r][c].  Thread = 8 columns, CTA = 2048 columns, grid.y = row slabs.
__global__ void __launch_bounds__(256) wide_head_bwd_kernel(const __nv_bfloat16* __restrict__ h, int64_t batch, int H,
                                                           const float* __restrict__ w_o, const float* __restrict__ dlog,
                                                           float inv_keep, __nv_bfloat16* __restrict__ dz,
                                                           float* __restrict__ part) {
    const int c8 = blockIdx.x * 256 + threadIdx.x;          // 8-column group
    if (c8 * 8 >= H) return;
    const int64_t rows_per = (batch + gridDim.y - 1) / gridDim.y;
    const int64_t r0 = int64_t(blockIdx.y) * rows_per, r1 = (r0 + rows_per < batch) ? r0 + rows_per : batch;
    float w0[8], w1[8], a0[8], a1[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        w0[i] = dz != nullptr ? __ldg(w_o + c8 * 8 + i) : 0.0f;
        w1[i] = dz != nullptr ? __ldg(w_o + H + c8 * 8 + i) : 0.0f;
        a0[i] = 0.0f;
        a1[i] = 0.0f;
    }
    for (int64_t rb = r0; rb < r1; rb += 4) {
        float2 dl[4];
        uint4 hv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {                      // 4 independent rows in flight
            const int64_t r = (rb + u < r1) ? rb + u : r1 - 1;
            dl[u] = __ldg(reinterpret_cast<const float2*>(dlog) + r);
            hv[u] = __ldg(reinterpret_cast<const uint4*>(h + r * H) + c8);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (rb + u >= r1) break;
            float f[8], o[8];
            unpack8(hv[u], f);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                a0[i] = fmaf(dl[u].x, f[i], a0[i]);
                a1[i] = fmaf(dl[u].y, f[i], a1[i]);
                const float dh = fmaf(dl[u].x, w0[i], dl[u].y * w1[i]);
                o[i] = f[i] > 0.0f ? dh * inv_keep : 0.0f;
            }
            uint4 ov;
            ov.x = pack_bf16(o[0], o[1]);
            ov.y = pack_bf16(o[2], o[3]);
            ov.z = pack_bf16(o[4], o[5]);
            ov.w = pack_bf16(o[6], o[7]);
            if (dz != nullptr) reinterpret_cast<uint4*>(dz + (rb + u) * H)[c8] = ov;
        }
    }
    float* p = part + size_t(blockIdx.y) * 2 * H;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        p[c8 * 8 + i] = a0[i];
        p[H + c8 * 8 + i] = a1[i];
    }
}

// per-slab column sums of a bf16 [batch, H] matrix (bias gradients)
__global__ void __launch_bounds__(256) wide_colsum_kernel(const __nv_bfloat16* __restrict__ dz, int64_t batch, int H,
                                                         float* __restrict__ part) {
    const int c8 = blockIdx.x * 256 + threadIdx.x;
    if (c8 * 8 >= H) return;
    const int64_t rows_per = (batch + gridDim.y - 1) / gridDim.y;
    const int64_t r0 = int64_t(blockIdx.y) * rows_per, r1 = (r0 + rows_per < batch) ? r0 + rows_per : batch;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int64_t rb = r0; rb < r1; rb += 4) {
        uint4 hv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t r = (rb + u < r1) ? rb + u : r1 - 1;
            hv[u] = __ldg(reinterpret_cast<const uint4*>(dz + r * H) + c8);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (rb + u >= r1) break;
            float f[8];
            unpack8(hv[u], f);
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] += f[i];
        }
    }
    float* p = part + size_t(blockIdx.y) * H;
#pragma unroll
    for (int i = 0; i < 8; ++i) p[c8 * 8 + i] = a[i];
}

// Tail tiles of a weight gradient (GemmArgs::tail_*): out += slab_1 + ... + slab_{n-1}, in that order, over the tiles
// [tail_first, tail_first + gridDim.y) only.  grid (tile_rows * tile_cols / 4 / 256 / 4, tail tiles), 4 float4 per thread.
__global__ void __launch_bounds__(256) wide_tail_fold_kernel(float* __restrict__ out, int64_t ldo, const float* __restrict__ slabs,
                                                            int64_t slab_stride, int n_slabs, int64_t tail_first, int tiles_n,
                                                            int tile_rows, int tile_cols) {
    const int64_t mn = tail_first + blockIdx.y;
    const int64_t m0 = (mn / tiles_n) * tile_rows, n0 = (mn % tiles_n) * tile_cols;
    const int c4 = tile_cols / 4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < tile_rows * c4; i += gridDim.x * blockDim.x) {
        const int64_t off = (m0 + i / c4) * ldo + n0 + (i % c4) * 4;
        float4 a = *reinterpret_cast<const float4*>(out + off);
        for (int s = 0; s < n_slabs; ++s) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(slabs + s * slab_stride + off));
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        *reinterpret_cast<float4*>(out + off) = a;
    }
}

// out[i] = sum over parts (in order) of part[p * stride + i]
__global__ void __launch_bounds__(256) wide_reduce_kernel(const float* __restrict__ part, int n_parts, int64_t stride,
                                                         int64_t n, float* __restrict__ out) {
    // CTA = 32 columns x 8 part groups: group g adds parts g, g + 8, ... in order, then the 8 group sums are added in group
    // order (fixed order; with 512 parts per column a thread per column was a 512-deep dependent chain on 16 CTAs)
    __shared__ float sh[8][32];
    const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
    const int64_t i = int64_t(blockIdx.x) * 32 + lane;
    float t = 0.0f;
    if (i < n)
        for (int p = grp; p < n_parts; p += 8) t += part[size_t(p) * stride + i];
    sh[grp][lane] = t;
    __syncthreads();
    if (grp == 0 && i < n) {
        float a = 0.0f;
#pragma unroll
        for (int k = 0; k < 8; ++k) a += sh[k][lane];
        out[i] = a;
    }
}

// head partials [n_parts][4] = (loss, db0, db1, -) -> grads[off_b_out .. +2] and grads[P]
__global__ void __launch_bounds__(256) wide_head_finish_kernel(const float* __restrict__ part, int n_parts,
                                                              float* __restrict__ db_out, float* __restrict__ loss_out,
                                                              float* __restrict__ correct_out) {
    __shared__ float sh[256][4];
    const int t = threadIdx.x;
    float s[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    for (int p = t; p < n_parts; p += 256) {               // fixed assignment + fixed tree: deterministic
        const float4 v = __ldg(reinterpret_cast<const float4*>(part) + p);
        s[0] += v.x; s[1] += v.y; s[2] += v.z; s[3] += v.w;
    }
    sh[t][0] = s[0]; sh[t][1] = s[1]; sh[t][2] = s[2]; sh[t][3] = s[3];
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (t < o) {
#pragma unroll
            for (int i = 0; i < 4; ++i) sh[t][i] += sh[t + o][i];
        }
        __syncthreads();
    }
    if (t == 0) {
        *loss_out = sh[0][0];
        if (db_out != nullptr) {
            db_out[0] = sh[0][1];
            db_out[1] = sh[0][2];
        }
        if (correct_out != nullptr) *correct_out = sh[0][3];
    }
}

// [splits][H, 64] fp32 -> [H, F] fp32: sums the split-K slabs of the first layer's weight gradient and drops the padding
__global__ void __launch_bounds__(256) unpad_kernel(const float* __restrict__ src, int64_t rows, int in, int pad,
                                                   int splits, float* __restrict__ dst) {
    const int64_t total = rows * in;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += int64_t(gridDim.x) * blockDim.x) {
        const int64_t r = i / in;
        float t = 0.0f;
        for (int s = 0; s < splits; ++s) t += src[(size_t(s) * rows + r) * pad + (i - r * in)];   // split-K slabs, in order
        dst[i] = t;
    }
}

// Adam over one parameter tensor; optionally re-emits its bf16 copy ([rows, pad] with `in` valid columns per row).
__global__ void __launch_bounds__(256) wide_adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ m, float* __restrict__ v, int64_t n, AdamS a,
                                                       __nv_bfloat16* __restrict__ wb, int in, int pad) {
    const int64_t stride = int64_t(gridDim.x) * blockDim.x;
    const int64_t t0 = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (in == pad && (n & 3) == 0) {
        const int64_t n4 = n >> 2;
        for (int64_t i = t0; i < n4; i += stride) {
            float4 pp = reinterpret_cast<float4*>(p)[i], mm = reinterpret_cast<float4*>(m)[i],
                   vv = reinterpret_cast<float4*>(v)[i];
            const float4 gg = __ldg(reinterpret_cast<const float4*>(g) + i);
            adam_one(pp.x, gg.x, mm.x, vv.x, a);
            adam_one(pp.y, gg.y, mm.y, vv.y, a);
            adam_one(pp.z, gg.z, mm.z, vv.z, a);
            adam_one(pp.w, gg.w, mm.w, vv.w, a);
            reinterpret_cast<float4*>(p)[i] = pp;
            reinterpret_cast<float4*>(m)[i] = mm;
            reinterpret_cast<float4*>(v)[i] = vv;
            if (wb != nullptr) {
                uint2 o;
                o.x = pack_bf16(pp.x, pp.y);
                o.y = pack_bf16(pp.z, pp.w);
                reinterpret_cast<uint2*>(wb)[i] = o;
            }
        }
    } else {
        for (int64_t i = t0; i < n; i += stride) {
            float pp = p[i], mm = m[i], vv = v[i];
            adam_one(pp, g[i], mm, vv, a);
            p[i] = pp;
            m[i] = mm;
            v[i] = vv;
            if (wb != nullptr) {
                const int64_t r = i / in;
                wb[r * pad + (i - r * in)] = __float2bfloat16_rn(pp);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// tensor maps and GEMM launches
// ------------------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_fn_t get_encode() {
    static encode_fn_t fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<encode_fn_t>(p);
    }
    return fn;
}

// bf16 row-major [rows, cols] (pitch in elements), 128-byte swizzle.
//   K-major operand (mn_groups == 0): 2-D map, box = 64 columns x box_rows rows.
//   MN-major operand: 3-D map (64 columns, all rows, cols/64 column groups), box = 64 x 64 x mn_groups -- one op drops
//   mn_groups consecutive 8 KiB atoms (64 k-lines of 64 M|N elements) into shared memory.
int make_map(CUtensorMap* tm, const void* base, int64_t rows, int64_t cols, int64_t pitch, int box_rows, int mn_groups = 0) {
    encode_fn_t enc = get_encode();
    if (enc == nullptr) {
        wx_set_error("cuTensorMapEncodeTiled is not available from this driver");
        return WX_ECUDA;
    }
    const cuuint64_t dims[3] = {mn_groups ? 64u : cuuint64_t(cols), cuuint64_t(rows), cuuint64_t(cols / 64)};
    const cuuint64_t strides[2] = {cuuint64_t(pitch) * 2, 128u};
    const cuuint32_t box[3] = {64u, cuuint32_t(box_rows), cuuint32_t(mn_groups)};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, mn_groups ? 3 : 2, const_cast<void*>(base), dims, strides,
                           box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        wx_set_error("cuTensorMapEncodeTiled failed with CUresult %d (base %p, %lld x %lld, pitch %lld)", int(r), base,
                     (long long)rows, (long long)cols, (long long)pitch);
        return WX_ECUDA;
    }
    return WX_OK;
}

template <bool A_K, bool B_K, int EPI>
int launch_gemm_t(wx_ctx* ctx, const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& g, cudaStream_t st) {
    static bool attr_set[64] = {};
    if (!attr_set[ctx->device & 63]) {
        WX_CUDA(cudaFuncSetAttribute(wide_gemm_kernel<A_K, B_K, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     int(GEMM_SMEM)));
        attr_set[ctx->device & 63] = true;
    }
    const int64_t tiles = ((g.M + BM - 1) / BM) * ((g.N + BN - 1) / BN) * g.split_k;
    const int64_t sms = ctx->num_sms - ctx->wide_reserved_sms;
    const int grid = int(tiles < sms ? tiles : sms);
    wide_gemm_kernel<A_K, B_K, EPI><<<grid, GEMM_THREADS, GEMM_SMEM, st>>>(ta, tb, g);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

// two-SM kernel: clusters of 2 CTAs, 256 x 256 (NB = 1) or 256 x 512 (NB = 2) tiles (wide_gemm2.cuh)
template <bool A_K, bool B_K, int EPI, int NB, int EW = 8>
int launch_gemm2_nb(wx_ctx* ctx, const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& g, cudaStream_t st) {
    static bool attr_set[64] = {};
    constexpr size_t GEMM2_SMEM = gemm2_smem(NB, EW);
    if (!attr_set[ctx->device & 63]) {
        WX_CUDA(cudaFuncSetAttribute(wide_gemm2_kernel<A_K, B_K, EPI, NB, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     int(GEMM2_SMEM)));
        attr_set[ctx->device & 63] = true;
    }
    const int64_t mn_tiles = (g.M / (2 * BM)) * (g.N / (NB * BN));
    const int64_t tiles = g.tail_split > 1 ? g.tail_first + (mn_tiles - g.tail_first) * g.tail_split : mn_tiles * g.split_k;
    const int64_t pairs = (ctx->num_sms - ctx->wide_reserved_sms) / 2;
    const int64_t clusters = tiles < pairs ? tiles : pairs;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(unsigned(clusters * 2));
    cfg.blockDim = dim3(gemm2_threads(EW));
    cfg.dynamicSmemBytes = GEMM2_SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    WX_CUDA(cudaLaunchKernelEx(&cfg, wide_gemm2_kernel<A_K, B_K, EPI, NB, EW>, ta, tb, g));
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

bool short_k_wide_epilogue() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("WX_WIDE_EW8");        // developer switch: keep 8 epilogue warps for the short-K product
        v = (e != nullptr && e[0] == '1') ? 0 : 1;
    }
    return v == 1;
}

int forced_tile_nb() {
    static int forced = -1;
    if (forced < 0) {
        const char* e = getenv("WX_WIDE_NB");         // developer switch for A/B timing of the two tile widths
        forced = (e != nullptr && (e[0] == '1' || e[0] == '2')) ? e[0] - '0' : 0;
    }
    return forced;
}

// Tile width of the two-SM kernel.  The 256 x 512 tile moves a quarter less data per FLOP through L2 and shared memory
// (the narrow tile asks the L2 for more than it delivers at the tensor pipe's full rate), but its accumulator takes all of
// TMEM, so a tile's epilogue is not hidden behind the next tile's MMAs -- which is why the epilogue warps do everything
// that does not need the accumulator (dropout bits, gate bits) while the MMAs run.  Wide when there are at least two
// wide tiles per SM pair.
int wide_tile_nb(const wx_ctx* ctx, const GemmArgs& g, int epi) {
    const bool fits = g.N % (2 * BN) == 0;
    if (g.nb) return (g.nb == 2 && fits) ? 2 : 1;
    if (forced_tile_nb()) return (forced_tile_nb() == 2 && fits) ? 2 : 1;
    const int64_t tiles2 = (g.M / (2 * BM)) * (g.N / (2 * BN)) * g.split_k, pairs = ctx->num_sms / 2;
    return (fits && tiles2 >= 2 * pairs) ? 2 : 1;
}

template <bool A_K, bool B_K, int EPI>
int launch_gemm2_t(wx_ctx* ctx, const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& g, cudaStream_t st) {
    // the first layer's forward product (K = one 64-deep block) is all epilogue: twice the epilogue warps
    if constexpr (EPI == EPI_FWD && A_K && B_K) {
        if (g.K <= 2 * BK && short_k_wide_epilogue()) return launch_gemm2_nb<true, true, EPI_FWD, 1, 16>(ctx, ta, tb, g, st);
    }
    if (wide_tile_nb(ctx, g, EPI) == 2) return launch_gemm2_nb<A_K, B_K, EPI, 2>(ctx, ta, tb, g, st);
    return launch_gemm2_nb<A_K, B_K, EPI, 1>(ctx, ta, tb, g, st);
}

bool two_sm_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("WX_WIDE_1CTA");      // developer switch for A/B timing of the two GEMM kernels
        v = (e != nullptr && e[0] == '1') ? 0 : 1;
    }
    return v == 1;
}

// D[M,N] = op(A) op(B): a_k / b_k say whether the operand is stored [M|N, K] (true) or [K, M|N] (false), row-major bf16.
int launch_gemm(wx_ctx* ctx, bool a_k, bool b_k, int epi, const void* A, int64_t lda, const void* B, int64_t ldb,
                GemmArgs g, cudaStream_t st) {
    CUtensorMap ta, tb;
    int rc;
    if (g.split_k < 1) g.split_k = 1;
    if (two_sm_enabled() && g.M % (2 * BM) == 0 && g.N % BN == 0 && (g.split_k == 1 || epi == EPI_DW || epi == EPI_PLAIN)) {
        // big aligned products: cluster pairs, each CTA stages its 128 rows of A and its 128-column half of B
        if (a_k) rc = make_map(&ta, A, g.M, g.K, lda, BM);
        else rc = make_map(&ta, A, g.K, g.M, lda, 64, BM / 64);
        if (rc != WX_OK) return rc;
        if (b_k) rc = make_map(&tb, B, g.N, g.K, ldb, BN / 2);
        else rc = make_map(&tb, B, g.K, g.N, ldb, 64, BN / 128);
        if (rc != WX_OK) return rc;
        g.err = reinterpret_cast<int*>(ctx->scratch) + 8;
        if (a_k && b_k) {
            if (epi == EPI_FWD) return launch_gemm2_t<true, true, EPI_FWD>(ctx, ta, tb, g, st);
            if (epi == EPI_PLAIN) return launch_gemm2_t<true, true, EPI_PLAIN>(ctx, ta, tb, g, st);
        } else if (a_k && !b_k) {
            if (epi == EPI_DX) return launch_gemm2_t<true, false, EPI_DX>(ctx, ta, tb, g, st);
            if (epi == EPI_PLAIN) return launch_gemm2_t<true, false, EPI_PLAIN>(ctx, ta, tb, g, st);
        } else if (!a_k && !b_k) {
            if (epi == EPI_DW) return launch_gemm2_t<false, false, EPI_DW>(ctx, ta, tb, g, st);
            if (epi == EPI_PLAIN) return launch_gemm2_t<false, false, EPI_PLAIN>(ctx, ta, tb, g, st);
        }
    }
    if (a_k) rc = make_map(&ta, A, g.M, g.K, lda, BM);
    else rc = make_map(&ta, A, g.K, g.M, lda, 64, BM / 64);
    if (rc != WX_OK) return rc;
    if (b_k) rc = make_map(&tb, B, g.N, g.K, ldb, BN);
    else rc = make_map(&tb, B, g.K, g.N, ldb, 64, BN / 64);
    if (rc != WX_OK) return rc;
    g.err = reinterpret_cast<int*>(ctx->scratch) + 8;   // sticky status word (wx_tc_last_status)
    if (g.split_k < 1) g.split_k = 1;
    if (a_k && b_k) {
        if (epi == EPI_FWD) return launch_gemm_t<true, true, EPI_FWD>(ctx, ta, tb, g, st);
        if (epi == EPI_PLAIN) return launch_gemm_t<true, true, EPI_PLAIN>(ctx, ta, tb, g, st);
    } else if (a_k && !b_k) {
        if (epi == EPI_DX) return launch_gemm_t<true, false, EPI_DX>(ctx, ta, tb, g, st);
        if (epi == EPI_PLAIN) return launch_gemm_t<true, false, EPI_PLAIN>(ctx, ta, tb, g, st);
    } else if (!a_k && !b_k) {
        if (epi == EPI_DW) return launch_gemm_t<false, false, EPI_DW>(ctx, ta, tb, g, st);
        if (epi == EPI_PLAIN) return launch_gemm_t<false, false, EPI_PLAIN>(ctx, ta, tb, g, st);
    }
    wx_set_error("wide gemm: unsupported layout/epilogue combination (a_k=%d b_k=%d epi=%d)", int(a_k), int(b_k), epi);
    return WX_EINVAL;
}

// How the weight-gradient product dW_layer = dZ^T H_below (hidden x hidden, K = batch) is cut into work units.
// (H/256)^2 tile pairs rarely fill the SM pairs evenly (256 tiles on 74 pairs = 3.46 rounds at H = 4096; 128 tiles of
// 256 x 512 = 1.73).  The whole rounds run as whole tiles; the tiles of the last, partly filled round are cut into 2 or 4
// ranges of the batch, range 0 accumulating into the gradient itself and ranges 1.. into workspace slabs that
// WX_WIDE_DW_FOLD adds in (a separate piece so it can run off the GEMM stream): 54 tiles x 4 ranges = 216 units = 2.92 short
// rounds instead of one long one.  Depends on the device's SM count only (NOT the reserved count: the DW and FOLD calls
// must agree).
struct DwPlan {
    int nb;               // tile width (1: 256 x 256, 2: 256 x 512); 0 = not the two-SM kernel's business
    int tail_split;       // 1 = no split
    int64_t tail_first;   // first split tile
    int64_t tail_tiles;   // number of split tiles
};

DwPlan plan_dw(const wx_ctx* ctx, const wx_wide_desc* d, int64_t batch, int layer) {
    DwPlan p = {0, 1, 0, 0};
    if (layer <= 1 || !two_sm_enabled()) return p;
    const int64_t H = d->hidden, pairs = ctx->num_sms / 2;
    p.nb = (H % (2 * BN) == 0 && forced_tile_nb() != 1) ? 2 : 1;
    if (forced_tile_nb() == 0 && (H / (2 * BM)) * (H / (2 * BN)) * 2 < pairs) p.nb = 1;
    const int64_t tiles = (H / (2 * BM)) * (H / (p.nb * BN));
    if (pairs < 1 || batch < 1024) return p;
    const int64_t rest = tiles % pairs;
    if (rest == 0) return p;
    double best = 1.0;
    for (int ts = 2; ts <= MAX_TAIL_SPLIT; ts *= 2) {
        if (batch / BK / ts < 8) break;                       // at least 8 K blocks per range
        const double rounds = double((rest * ts + pairs - 1) / pairs) / ts;
        if (rounds < best - 1e-9) {
            best = rounds;
            p.tail_split = ts;
        }
    }
    if (p.tail_split > 1) {
        p.tail_first = tiles - rest;
        p.tail_tiles = rest;
    }
    return p;
}

AdamS adam_consts(const wx_wide_desc* d, int64_t step) {
    AdamS a;
    const double bc1 = 1.0 - pow(d->beta1, double(step)), bc2 = 1.0 - pow(d->beta2, double(step));
    a.w_lerp = float(1.0 - d->beta1);
    a.beta2 = float(d->beta2);
    a.one_m_beta2 = float(1.0 - d->beta2);
    a.neg_step = float(-(d->lr / bc1));
    a.bc2_sqrt = float(sqrt(bc2));
    a.eps = float(d->eps);
    return a;
}

int grid_for(const wx_ctx* ctx, int64_t n, int per_thread = 1) {
    int64_t b = (n / per_thread + 255) / 256;
    const int64_t cap = int64_t(ctx->num_sms) * 8;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return int(b);
}

}  // namespace

extern "C" {

int64_t wx_wide_n_params(const wx_wide_desc* desc) {
    if (check_desc(desc, "wx_wide_n_params") != WX_OK) return -1;
    Layout L;
    make_layout(desc, 128, &L);
    return L.P;
}

int64_t wx_wide_workspace_bytes(const wx_wide_desc* desc, int64_t batch) {
    if (check_desc(desc, "wx_wide_workspace_bytes") != WX_OK) return -1;
    if (batch < 128 || batch % 128 != 0) {
        wx_set_error("wx_wide_workspace_bytes: batch=%lld must be a positive multiple of 128", (long long)batch);
        return -1;
    }
    Layout L;
    make_layout(desc, batch, &L);
    return int64_t(L.total);
}

/* Self-test / building block: D[M,N] (fp32, pitch N) = op(A) op(B) (+ bias), bf16 operands, fp32 accumulate. */
int wx_wide_gemm(wx_ctx* ctx, int a_kmajor, int b_kmajor, const void* A, int64_t lda, const void* B, int64_t ldb,
                 int64_t M, int64_t N, int64_t K, const float* bias, float* out, void* stream) {
    WX_CHECK_ARG(ctx && A && B && out, "wx_wide_gemm: NULL argument");
    WX_CHECK_ARG(M >= 1 && N >= 16 && N % 16 == 0 && K >= 64 && K % 64 == 0, "wx_wide_gemm: need N %% 16 == 0, K %% 64 == 0");
    WX_CHECK_ARG((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(out)) % 16 == 0 &&
                     lda % 8 == 0 && ldb % 8 == 0,
                 "wx_wide_gemm: operands must be 16-byte aligned with pitches that are multiples of 8 elements");
    WX_CHECK_ARG((a_kmajor || M % 64 == 0) && (b_kmajor || N % 64 == 0),
                 "wx_wide_gemm: an operand stored [K, M|N] needs M|N %% 64 == 0");
    WX_CUDA(cudaSetDevice(ctx->device));
    GemmArgs g = {};
    g.M = M; g.N = N; g.K = K;
    g.out = out; g.ldo = N; g.bias = bias;
    g.inv_keep = 1.0f; g.keep_thresh16 = 65536u;
    return launch_gemm(ctx, a_kmajor != 0, b_kmajor != 0, EPI_PLAIN, A, lda, B, ldb, g, (cudaStream_t)stream);
}

int wx_wide_cast_weights(wx_ctx* ctx, const wx_wide_desc* desc, const float* params, void* ws, void* stream) {
    WX_CHECK_ARG(ctx && params && ws, "wx_wide_cast_weights: NULL argument");
    int rc = check_desc(desc, "wx_wide_cast_weights");
    if (rc != WX_OK) return rc;
    WX_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 1023) == 0, "wx_wide_cast_weights: ws must be 1024-byte aligned");
    WX_CUDA(cudaSetDevice(ctx->device));
    Layout L;
    make_layout(desc, 128, &L);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* w = static_cast<unsigned char*>(ws);
    const int64_t H = desc->hidden;
    cast_pad_kernel<<<grid_for(ctx, H * KPAD), 256, 0, st>>>(params + L.off_w[1], H, desc->n_features, KPAD,
                                                            reinterpret_cast<__nv_bfloat16*>(w + L.wb[1]));
    WX_LAUNCH_CHECK(ctx);
    for (int l = 2; l <= desc->n_hidden; ++l) {
        cast_kernel<<<grid_for(ctx, H * H, 4), 256, 0, st>>>(params + L.off_w[l], H * H,
                                                           reinterpret_cast<__nv_bfloat16*>(w + L.wb[l]));
        WX_LAUNCH_CHECK(ctx);
    }
    return WX_OK;
}

}  // extern "C"

namespace {

// x -> X0 -> H_1 .. H_L (bf16) in the workspace: the hidden stack shared by the training forward and the validation pass
int hidden_forward(wx_ctx* ctx, const wx_wide_desc* desc, const Layout& L, const float* params, unsigned char* w,
                   const float* x, int64_t batch, bool drop, uint64_t seed, uint64_t step, cudaStream_t st) {
    const int64_t H = desc->hidden;
    const float keep_p = 1.0f - desc->dropout_p;
    __nv_bfloat16* x0 = reinterpret_cast<__nv_bfloat16*>(w + L.x0);
    cast_pad_kernel<<<grid_for(ctx, batch * KPAD), 256, 0, st>>>(x, batch, desc->n_features, KPAD, x0);
    WX_LAUNCH_CHECK(ctx);
    for (int l = 1; l <= desc->n_hidden; ++l) {
        GemmArgs g = {};
        g.M = batch; g.N = H; g.K = (l == 1) ? KPAD : H;
        g.out = w + L.h[l]; g.ldo = H;
        g.bias = params + L.off_b[l];
        g.inv_keep = drop ? 1.0f / keep_p : 1.0f;
        g.keep_thresh16 = drop ? uint32_t(lrintf(keep_p * 65536.0f)) : 65536u;
        g.seed_lo = uint32_t(seed); g.seed_hi = uint32_t(seed >> 32);
        g.step = uint32_t(step); g.layer = uint32_t(l);
        const void* a = (l == 1) ? static_cast<const void*>(x0) : static_cast<const void*>(w + L.h[l - 1]);
        const int rc = launch_gemm(ctx, true, true, EPI_FWD, a, g.K, w + L.wb[l], g.K, g, st);
        if (rc != WX_OK) return rc;
    }
    return WX_OK;
}

}  // namespace

extern "C" {

/* Forward through the loss, plus the head's gradients and dZ of the last hidden layer.
 *   x      fp32 [batch, F] row-major, standardised (the output of wx_normalize_pack)
 *   grads  fp32 [P+1]: receives db_o and, at [P], sum_rows(CE) * grad_scale
 *   grad_scale = 1 / (world * batch): gradients and loss come out pre-scaled for a SUM all-reduce (DDP, :136)
 *   dropout_mode: WX_DROPOUT_OFF or WX_DROPOUT_PHILOX (keep <=> 16 Philox bits < round((1-p) * 65536);
 *                 counter (row, col/8, layer, step), key seed -- oracle/wide_ref.py keep_mask) */
int wx_wide_forward(wx_ctx* ctx, const wx_wide_desc* desc, const float* params, void* ws, const float* x,
                    const int64_t* y, int64_t batch, int dropout_mode, uint64_t seed, uint64_t step, float grad_scale,
                    float* grads, void* stream) {
    WX_CHECK_ARG(ctx && params && ws && x && y && grads, "wx_wide_forward: NULL argument");
    int rc = check_desc(desc, "wx_wide_forward");
    if (rc != WX_OK) return rc;
    WX_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 1023) == 0, "wx_wide_forward: ws must be 1024-byte aligned");
    WX_CHECK_ARG(batch >= 128 && batch % 128 == 0, "wx_wide_forward: batch=%lld must be a positive multiple of 128",
                 (long long)batch);
    WX_CHECK_ARG(dropout_mode == WX_DROPOUT_OFF || dropout_mode == WX_DROPOUT_PHILOX,
                 "wx_wide_forward: dropout_mode=%d (OFF or PHILOX)", dropout_mode);
    WX_CUDA(cudaSetDevice(ctx->device));
    Layout L;
    make_layout(desc, batch, &L);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* w = static_cast<unsigned char*>(ws);
    const int64_t H = desc->hidden;
    const int nl = desc->n_hidden;
    const bool drop = dropout_mode == WX_DROPOUT_PHILOX && desc->dropout_p > 0.0f;
    const float keep_p = 1.0f - desc->dropout_p;
    rc = hidden_forward(ctx, desc, L, params, w, x, batch, drop, seed, step, st);
    if (rc != WX_OK) return rc;
    const __nv_bfloat16* hl = reinterpret_cast<const __nv_bfloat16*>(w + L.h[nl]);
    float* dlog = reinterpret_cast<float*>(w + L.dlog);
    float* part = reinterpret_cast<float*>(w + L.part);
    float* head_part = part + size_t(HEAD_SLABS) * 2 * H;       // [grid][4]
    // logits, loss, dlogits AND dZ_L in one pass over H_L; dW_o (the other reader of H_L) is WX_WIDE_HEAD_DW of
    // wx_wide_backward_layer(layer = n_hidden), so the caller can keep it off the stream the GEMMs are waiting on
    const int hgrid = launch_head(ctx, hl, batch, int(H), params + L.off_w[nl + 1], params + L.off_b[nl + 1], y, grad_scale,
                                  dlog, head_part, drop ? 1.0f / keep_p : 1.0f, reinterpret_cast<__nv_bfloat16*>(w + L.dz[0]),
                                  st);
    WX_LAUNCH_CHECK(ctx);
    wide_head_finish_kernel<<<1, 256, 0, st>>>(head_part, hgrid, grads + L.off_b[nl + 1], grads + L.P, nullptr);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

/* Backward of hidden layer `layer` (call with layer = n_hidden, ..., 1 after wx_wide_forward): writes dW_layer and
 * db_layer into grads and, for layer > 1, dZ of the layer below into the workspace. */
int wx_wide_backward_layer(wx_ctx* ctx, const wx_wide_desc* desc, void* ws, int64_t batch, int layer, int dropout_mode,
                           int parts, float* grads, void* stream) {
    WX_CHECK_ARG(ctx && ws && grads, "wx_wide_backward_layer: NULL argument");
    int rc = check_desc(desc, "wx_wide_backward_layer");
    if (rc != WX_OK) return rc;
    WX_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 1023) == 0, "wx_wide_backward_layer: ws must be 1024-byte aligned");
    WX_CHECK_ARG(batch >= 128 && batch % 128 == 0, "wx_wide_backward_layer: batch=%lld must be a positive multiple of 128",
                 (long long)batch);
    const int nl = desc->n_hidden;
    WX_CHECK_ARG(layer >= 1 && layer <= nl, "wx_wide_backward_layer: layer=%d outside 1..%d", layer, nl);
    WX_CHECK_ARG(parts >= 1 && parts <= 63, "wx_wide_backward_layer: parts=%d (bit 0 dW, bit 1 db, bit 2 dZ below, bit 3 fold, bit 4 head dW, bit 5 db from the fused partials)", parts);
    WX_CUDA(cudaSetDevice(ctx->device));
    Layout L;
    make_layout(desc, batch, &L);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* w = static_cast<unsigned char*>(ws);
    const int64_t H = desc->hidden;
    const bool drop = dropout_mode == WX_DROPOUT_PHILOX && desc->dropout_p > 0.0f;
    const void* dz = w + L.dz[(nl - layer) & 1];
    void* dz_below = w + L.dz[(nl - layer + 1) & 1];
    const void* h_below = (layer == 1) ? static_cast<const void*>(w + L.x0) : static_cast<const void*>(w + L.h[layer - 1]);
    const int64_t in = (layer == 1) ? KPAD : H;
    // head: dW_o[k][c] = sum_r dlogits[r][k] H_L[r][c] (per-slab partial sums, then an ordered reduce)
    if ((parts & WX_WIDE_HEAD_DW) && layer == nl) {
        float* part = reinterpret_cast<float*>(w + L.part2);
        int slabs = int(batch / 128 < HEAD_SLABS ? batch / 128 : HEAD_SLABS);
        dim3 bg((unsigned)((H / 8 + 255) / 256), (unsigned)slabs);
        wide_head_bwd_kernel<<<bg, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(w + L.h[nl]), batch, int(H),
                                                nullptr, reinterpret_cast<const float*>(w + L.dlog), 1.0f,
                                                nullptr, part);
        WX_LAUNCH_CHECK(ctx);
        wide_reduce_kernel<<<int((2 * H + 31) / 32), 256, 0, st>>>(part, slabs, 2 * H, 2 * H, grads + L.off_w[nl + 1]);
        WX_LAUNCH_CHECK(ctx);
    }
    // dW = dZ^T H_below : A = dZ stored [batch, H] = [K, M], B = H_below stored [batch, in] = [K, N]
    if (parts & WX_WIDE_DW) {
        GemmArgs g = {};
        g.M = H; g.N = in; g.K = batch;
        g.inv_keep = 1.0f; g.keep_thresh16 = 65536u;
        int splits = 1;
        if (layer == 1) {
            // 32 output tiles cannot fill 148 SMs: cut the batch into split-K ranges, one fp32 slab each (summed by unpad)
            const int64_t tiles = ((H + BM - 1) / BM);
            while (splits < MAX_SPLIT_K && tiles * splits * 2 <= ctx->num_sms && batch / 64 / (splits * 2) >= 4) splits *= 2;
            g.split_k = splits;
            g.split_stride = H * KPAD;
            g.out = w + L.dw1tmp; g.ldo = KPAD;
        } else {
            g.out = grads + L.off_w[layer]; g.ldo = H;
            const DwPlan pl = plan_dw(ctx, desc, batch, layer);
            g.nb = pl.nb;
            if (pl.tail_split > 1) {
                g.tail_split = pl.tail_split;
                g.tail_first = pl.tail_first;
                g.tail_off = reinterpret_cast<float*>(w + L.dwtmp[layer]) - (grads + L.off_w[layer]);
                g.tail_stride = H * H;
            }
        }
        rc = launch_gemm(ctx, false, false, EPI_DW, dz, H, h_below, in, g, st);
        if (rc != WX_OK) return rc;
        if (layer == 1) {
            unpad_kernel<<<grid_for(ctx, H * desc->n_features), 256, 0, st>>>(
                reinterpret_cast<const float*>(w + L.dw1tmp), H, desc->n_features, KPAD, splits, grads + L.off_w[1]);
            WX_LAUNCH_CHECK(ctx);
        }
    }
    if (parts & WX_WIDE_DW_FOLD) {
        const DwPlan pl = plan_dw(ctx, desc, batch, layer);
        if (pl.tail_split > 1) {
            const int tile_cols = pl.nb * BN;
            dim3 fg(unsigned(2 * BM * tile_cols / 4 / 256 / 4), unsigned(pl.tail_tiles));
            wide_tail_fold_kernel<<<fg, 256, 0, st>>>(grads + L.off_w[layer], H, reinterpret_cast<const float*>(w + L.dwtmp[layer]),
                                                     H * H, pl.tail_split - 1, pl.tail_first, int(H / tile_cols), 2 * BM,
                                                     tile_cols);
            WX_LAUNCH_CHECK(ctx);
        }
    }
    // db = column sums of dZ
    if (parts & WX_WIDE_DB) {
        float* part = reinterpret_cast<float*>(w + L.part);
        int slabs = int(batch / 128 < COLSUM_SLABS ? batch / 128 : COLSUM_SLABS);
        dim3 cg((unsigned)((H / 8 + 255) / 256), (unsigned)slabs);
        wide_colsum_kernel<<<cg, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(dz), batch, int(H), part);
        WX_LAUNCH_CHECK(ctx);
        wide_reduce_kernel<<<int((H + 31) / 32), 256, 0, st>>>(part, slabs, H, H, grads + L.off_b[layer]);
        WX_LAUNCH_CHECK(ctx);
    }
    // db from the 32-row partial column sums that the dZ product of the layer above left behind (no pass over dZ)
    if (parts & WX_WIDE_DB_FUSED) {
        WX_CHECK_ARG(layer < nl, "wx_wide_backward_layer: WX_WIDE_DB_FUSED needs a layer below the top hidden layer "
                     "(its dZ comes from the head kernel, not from a dZ product); use WX_WIDE_DB there");
        wide_reduce_kernel<<<int((H + 31) / 32), 256, 0, st>>>(reinterpret_cast<const float*>(w + L.cpart[layer]),
                                                                int((batch + 31) / 32), H, H, grads + L.off_b[layer]);
        WX_LAUNCH_CHECK(ctx);
    }
    // dZ_below = (dZ W) * gate(H_below): A = dZ [batch, H] K-major, B = W stored [H(out), H(in)] = [K, N]
    if (layer > 1 && (parts & WX_WIDE_DZ)) {
        GemmArgs g = {};
        g.M = batch; g.N = H; g.K = H;
        g.out = dz_below; g.ldo = H;
        g.gate = reinterpret_cast<const __nv_bfloat16*>(h_below); g.ldg = H;
        g.inv_keep = drop ? 1.0f / (1.0f - desc->dropout_p) : 1.0f;
        g.keep_thresh16 = 65536u;
        g.colsum_part = reinterpret_cast<float*>(w + L.cpart[layer - 1]);   // db of the layer below, as 32-row partials
        rc = launch_gemm(ctx, true, false, EPI_DX, dz, H, w + L.wb[layer], H, g, st);
        if (rc != WX_OK) return rc;
    }
    return WX_OK;
}

/* Adam for one layer's tensors (layer = 1..n_hidden, n_hidden+1 = the head): lets the caller run the update of a layer
 * on a side stream as soon as its gradient slice is final, under the backward GEMMs of the layers below. */
int wx_wide_adam_layer(wx_ctx* ctx, const wx_wide_desc* desc, float* params, const float* grads, float* m, float* v,
                       int64_t step, void* ws, int layer, void* stream) {
    WX_CHECK_ARG(ctx && params && grads && m && v && ws, "wx_wide_adam_layer: NULL argument");
    int rc = check_desc(desc, "wx_wide_adam_layer");
    if (rc != WX_OK) return rc;
    WX_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 1023) == 0, "wx_wide_adam_layer: ws must be 1024-byte aligned");
    WX_CHECK_ARG(step >= 1, "wx_wide_adam_layer: step=%lld must be >= 1", (long long)step);
    const int nl = desc->n_hidden;
    WX_CHECK_ARG(layer >= 1 && layer <= nl + 1, "wx_wide_adam_layer: layer=%d outside 1..%d (%d = head)", layer, nl + 1, nl + 1);
    WX_CUDA(cudaSetDevice(ctx->device));
    Layout L;
    make_layout(desc, 128, &L);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* w = static_cast<unsigned char*>(ws);
    const AdamS a = adam_consts(desc, step);
    const int64_t H = desc->hidden;
    const int64_t in = (layer == 1) ? desc->n_features : H;
    const int64_t rows = (layer <= nl) ? H : desc->classes;
    const int64_t nw = rows * in, w0 = L.off_w[layer], b0 = L.off_b[layer];
    __nv_bfloat16* wb = (layer <= nl) ? reinterpret_cast<__nv_bfloat16*>(w + L.wb[layer]) : nullptr;
    wide_adam_kernel<<<grid_for(ctx, nw, 4), 256, 0, st>>>(params + w0, grads + w0, m + w0, v + w0, nw, a, wb,
                                                         wb ? int(in) : 1, wb ? (layer == 1 ? KPAD : int(in)) : 1);
    WX_LAUNCH_CHECK(ctx);
    wide_adam_kernel<<<grid_for(ctx, rows), 256, 0, st>>>(params + b0, grads + b0, m + b0, v + b0, rows, a, nullptr, 1, 1);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

/* Adam (train_lightning_ddp.py:88 arithmetic) over all P parameters; the hidden layers' bf16 copies in the workspace are
 * rewritten in the same pass.  step is 1-based. */
int wx_wide_adam(wx_ctx* ctx, const wx_wide_desc* desc, float* params, const float* grads, float* m, float* v,
                 int64_t step, void* ws, void* stream) {
    int rc = check_desc(desc, "wx_wide_adam");
    if (rc != WX_OK) return rc;
    for (int l = 1; l <= desc->n_hidden + 1; ++l) {
        rc = wx_wide_adam_layer(ctx, desc, params, grads, m, v, step, ws, l, stream);
        if (rc != WX_OK) return rc;
    }
    return WX_OK;
}

/* Validation pass (train_lightning_ddp.py:73-85 for the wide net): eval-mode forward (no dropout), out[0] = sum of the
 * per-row cross-entropy, out[1] = number of rows whose argmax equals the label (fp32 device scalars). */
int wx_wide_eval(wx_ctx* ctx, const wx_wide_desc* desc, const float* params, void* ws, const float* x, const int64_t* y,
                 int64_t batch, float* out, void* stream) {
    WX_CHECK_ARG(ctx && params && ws && x && y && out, "wx_wide_eval: NULL argument");
    int rc = check_desc(desc, "wx_wide_eval");
    if (rc != WX_OK) return rc;
    WX_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 1023) == 0, "wx_wide_eval: ws must be 1024-byte aligned");
    WX_CHECK_ARG(batch >= 128 && batch % 128 == 0, "wx_wide_eval: batch=%lld must be a positive multiple of 128",
                 (long long)batch);
    WX_CUDA(cudaSetDevice(ctx->device));
    Layout L;
    make_layout(desc, batch, &L);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* w = static_cast<unsigned char*>(ws);
    const int nl = desc->n_hidden;
    rc = hidden_forward(ctx, desc, L, params, w, x, batch, false, 0, 0, st);
    if (rc != WX_OK) return rc;
    float* part = reinterpret_cast<float*>(w + L.part);
    float* head_part = part + size_t(HEAD_SLABS) * 2 * desc->hidden;
    const int hgrid = launch_head(ctx, reinterpret_cast<const __nv_bfloat16*>(w + L.h[nl]), batch, desc->hidden,
                                  params + L.off_w[nl + 1], params + L.off_b[nl + 1], y, 1.0f, nullptr, head_part, 1.0f,
                                  nullptr, st);
    WX_LAUNCH_CHECK(ctx);
    wide_head_finish_kernel<<<1, 256, 0, st>>>(head_part, hgrid, nullptr, out, out + 1);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

/* Leave `n` SMs (0..64, even) out of the wide GEMM grids so a collective running on another stream (the per-layer gradient
 * all-reduce) always has somewhere to run instead of queueing behind one-CTA-per-SM persistent kernels. */
int wx_wide_reserve_sms(wx_ctx* ctx, int n) {
    WX_CHECK_ARG(ctx != nullptr, "wx_wide_reserve_sms: ctx is NULL");
    WX_CHECK_ARG(n >= 0 && n <= 64 && n % 2 == 0 && n < ctx->num_sms, "wx_wide_reserve_sms: n=%d must be even, in 0..64", n);
    ctx->wide_reserved_sms = n;
    return WX_OK;
}

}  // extern "C"
