// Device-side building blocks of the fused weather-MLP step (K2/K3/K5), shared by the grid-wide step
// kernel and the single-CTA multi-step kernel.
//
// Mapping (fp32 FFMA path, strict parity with torch's fp32 arithmetic):
//   * one warp walks chunks of R = 8 rows; lane l owns hidden units j = l and j = l + 32:
//     their W1 rows, b1, W2 columns live in registers for the whole kernel, and so do the
//     running gradient sums for exactly those parameters -- the batch reduction of dW1/dW2/db1
//     costs no shuffles and no atomics;
//   * x rows are staged global -> shared with cp.async (double buffered per warp) and read back
//     as warp-broadcast LDS.128;
//   * the only cross-lane traffic is the 2-logit reduction: the 16 partial logits of a chunk are
//     reduced with one transposing butterfly (15+1 shuffles per 8 rows), after which lane l holds
//     logit[row l>>2][class (l>>1)&1]; softmax / loss / dlogits are then computed SIMD across rows.
#pragma once

#include "common.cuh"

namespace wxk {

constexpr int R = 8;   // rows per warp chunk

enum Loader { LOAD_ROWMAJOR = 0, LOAD_RAW_F32 = 1, LOAD_RAW_F64 = 2 };

template <int F>
struct Dim {
    static constexpr int FS = (F + 3) & ~3;                 // smem row stride in floats (16 B aligned rows)
    static constexpr int P = 64 * F + 64 + 128 + 2;
    static constexpr int OFF_B1 = 64 * F;
    static constexpr int OFF_W2 = 64 * F + 64;
    static constexpr int OFF_B2 = 64 * F + 64 + 128;
    static constexpr int SLAB = (P + 1 + 3) & ~3;           // grads + loss, padded to 16 B
};

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11).  Dropout stream contract (WX_DROPOUT_PHILOX; specified in oracle/philox_ref.py):
// one 64-bit keep mask per batch position b (bit j = keep hidden unit j) from TWO Philox calls q = 0, 1,
//   key = (seed_lo, seed_hi), counter = (b_lo, b_hi, step_lo, (step_hi << 3) | q) -> words W[4q .. 4q+3];
// word W[k] serves units 8k .. 8k+7 through its first eight base-5 digits:
//   s_0 = W[k]; digit_t = hi32(5 * s_t); s_{t+1} = lo32(5 * s_t);   keep_{8k+t} <=> digit_t != 0
// so P(keep) = 1 - ceil(2^32 / 5) / 2^32 = 0.8 (nn.Dropout(0.2), train_lightning_ddp.py:60) at one 32x32->64 multiply
// per unit.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0;
        k.y += W1;
    }
    return c;
}

// one base-5 digit: returns digit != 0 (keep) and advances the state
__device__ __forceinline__ bool keep_digit(uint32_t& s) {
    const uint64_t prod = uint64_t(s) * 5ull;
    s = uint32_t(prod);
    return uint32_t(prod >> 32) != 0u;
}

// keep bits of units 8k .. 8k+7 (bit t = unit 8k+t) from word W[k]
__device__ __forceinline__ uint32_t keep8_from_word(uint32_t s) {
    uint32_t bits = 0;
#pragma unroll
    for (int t = 0; t < 8; ++t) bits |= uint32_t(keep_digit(s)) << t;
    return bits;
}

// the eight words of batch position b (two Philox calls)
__device__ __forceinline__ void philox_keep_words(uint64_t b, uint2 key, uint64_t step, uint32_t (&w)[8]) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const uint4 o = philox4x32_10(make_uint4(uint32_t(b), uint32_t(b >> 32), uint32_t(step),
                                                 (uint32_t(step >> 32) << 3) | uint32_t(q)), key);
        w[4 * q] = o.x; w[4 * q + 1] = o.y; w[4 * q + 2] = o.z; w[4 * q + 3] = o.w;
    }
}

// keep bits of units 32q .. 32q+31 of batch position b: Philox call q
__device__ __forceinline__ uint32_t philox_keep_half(uint64_t b, int q, uint2 key, uint64_t step) {
    const uint4 o = philox4x32_10(make_uint4(uint32_t(b), uint32_t(b >> 32), uint32_t(step),
                                             (uint32_t(step >> 32) << 3) | uint32_t(q)), key);
    return keep8_from_word(o.x) | (keep8_from_word(o.y) << 8) | (keep8_from_word(o.z) << 16) | (keep8_from_word(o.w) << 24);
}

// the whole 64-bit keep mask of batch position b; used where one thread owns a row
__device__ __forceinline__ uint64_t philox_keep_mask(uint64_t b, uint2 key, uint64_t step) {
    return uint64_t(philox_keep_half(b, 0, key, step)) | (uint64_t(philox_keep_half(b, 1, key, step)) << 32);
}

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// Per-lane model slice + gradient accumulators.
template <int F>
struct LaneState {
    float w1a[F], w1b[F];          // W1[lane][*], W1[lane+32][*]
    float b1a, b1b;
    float w2a0, w2b0, w2a1, w2b1;  // W2[0][lane], W2[0][lane+32], W2[1][lane], W2[1][lane+32]
    float b2_0, b2_1;
    // gradient sums (unscaled by anything but the per-row `scale` already folded into dlogits)
    float gw1a[F], gw1b[F];
    float gb1a, gb1b;
    float gw2a0, gw2b0, gw2a1, gw2b1;
    float gb2_0, gb2_1;            // per-lane partial (only lanes with (lane&3)==0 add)
    float loss;                    // per-lane partial
    int correct;                   // eval only

    __device__ __forceinline__ void load_weights(const float* __restrict__ p, int lane) {
        using D = Dim<F>;
#pragma unroll
        for (int f = 0; f < F; ++f) {
            w1a[f] = p[lane * F + f];
            w1b[f] = p[(lane + 32) * F + f];
        }
        b1a = p[D::OFF_B1 + lane];
        b1b = p[D::OFF_B1 + lane + 32];
        w2a0 = p[D::OFF_W2 + lane];
        w2b0 = p[D::OFF_W2 + lane + 32];
        w2a1 = p[D::OFF_W2 + 64 + lane];
        w2b1 = p[D::OFF_W2 + 64 + lane + 32];
        b2_0 = p[D::OFF_B2];
        b2_1 = p[D::OFF_B2 + 1];
    }
    __device__ __forceinline__ void zero_grads() {
#pragma unroll
        for (int f = 0; f < F; ++f) gw1a[f] = gw1b[f] = 0.0f;
        gb1a = gb1b = gw2a0 = gw2b0 = gw2a1 = gw2b1 = gb2_0 = gb2_1 = loss = 0.0f;
        correct = 0;
    }
    // Add this warp's sums into a CTA slab laid out like the flat parameter vector (+ loss at [P]).
    // first == true overwrites.  Each (lane, element) pair is unique, so no intra-warp conflicts.
    __device__ __forceinline__ void flush_to_slab(float* slab, int lane, bool first) {
        using D = Dim<F>;
        const float gb2_0s = warp_sum(gb2_0), gb2_1s = warp_sum(gb2_1), ls = warp_sum(loss);
#define WX_ACC(idx, val)                \
    do {                                \
        if (first)                      \
            slab[idx] = (val);          \
        else                            \
            slab[idx] += (val);         \
    } while (0)
#pragma unroll
        for (int f = 0; f < F; ++f) {
            WX_ACC(lane * F + f, gw1a[f]);
            WX_ACC((lane + 32) * F + f, gw1b[f]);
        }
        WX_ACC(D::OFF_B1 + lane, gb1a);
        WX_ACC(D::OFF_B1 + lane + 32, gb1b);
        WX_ACC(D::OFF_W2 + lane, gw2a0);
        WX_ACC(D::OFF_W2 + lane + 32, gw2b0);
        WX_ACC(D::OFF_W2 + 64 + lane, gw2a1);
        WX_ACC(D::OFF_W2 + 64 + lane + 32, gw2b1);
        if (lane == 0) {
            WX_ACC(D::OFF_B2, gb2_0s);
            WX_ACC(D::OFF_B2 + 1, gb2_1s);
            WX_ACC(D::P, ls);
        }
#undef WX_ACC
    }
};

// Per-chunk row metadata held in registers of lanes 0..R-1.
struct ChunkMeta {
    int64_t rowid;      // global row of chunk row `lane` (lane < R), -1 if past the batch end
    int y;              // label of that row
    uint32_t mlo, mhi;  // dropout keep bits of that row (WX_DROPOUT_MASK)
};

// Transposing butterfly: in: v[2r+c] = this lane's partial of logit c of row r (16 values);
// out: total over the warp of logit[(lane>>2)][(lane>>1)&1].
__device__ __forceinline__ float reduce16_transpose(const float (&v)[16], int lane) {
    float u[8], t[4], s[2];
    {
        const bool hi = lane & 16;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float send = hi ? v[i] : v[i + 8];
            const float keep = hi ? v[i + 8] : v[i];
            u[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool hi = lane & 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float send = hi ? u[i] : u[i + 4];
            const float keep = hi ? u[i + 4] : u[i];
            t[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool hi = lane & 4;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const float send = hi ? t[i] : t[i + 2];
            const float keep = hi ? t[i + 2] : t[i];
            s[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
    }
    const bool hi = lane & 2;
    const float send = hi ? s[0] : s[1];
    const float keep = hi ? s[1] : s[0];
    float r = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    r += __shfl_xor_sync(0xffffffffu, r, 1);
    return r;
}

// keep-bit word for a chunk: bit 2r = keep unit `lane` of chunk row r, bit 2r+1 = keep unit lane+32.
__device__ __forceinline__ uint32_t chunk_keep_bits(int mode, const ChunkMeta& cm, int lane, int64_t chunk_pos0,
                                                    uint2 key, uint32_t step_lo, uint32_t step_hi) {
    if (mode == WX_DROPOUT_OFF) return 0xffffu;
    uint32_t kb = 0;
    if (mode == WX_DROPOUT_MASK) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint32_t lo = __shfl_sync(0xffffffffu, cm.mlo, r);
            const uint32_t hi = __shfl_sync(0xffffffffu, cm.mhi, r);
            kb |= ((lo >> lane) & 1u) << (2 * r);
            kb |= ((hi >> lane) & 1u) << (2 * r + 1);
        }
        return kb;
    }
    // WX_DROPOUT_PHILOX: lane -> (row r = lane >> 2 of the chunk, Philox call q = lane & 1 of that row): lane 4r holds the
    // keep bits of units 0..31 of row r, lane 4r+1 those of units 32..63; every lane then picks its two units' bits.
    const uint64_t b = uint64_t(chunk_pos0) + uint64_t(lane >> 2);
    const uint32_t half = philox_keep_half(b, lane & 1, key, (uint64_t(step_hi) << 32) | step_lo);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const uint32_t lo = __shfl_sync(0xffffffffu, half, 4 * r);
        const uint32_t hi = __shfl_sync(0xffffffffu, half, 4 * r + 1);
        kb |= ((lo >> lane) & 1u) << (2 * r);
        kb |= ((hi >> lane) & 1u) << (2 * r + 1);
    }
    return kb;
}

// One chunk: forward for R rows, softmax/loss, and (TRAIN) backward accumulation.
//   xs: smem [R][FS] fp32 standardised features of the chunk (rows past the end hold zeros)
//   MODE: 0 = train (grads), 1 = eval (loss sum + correct), 2 = predict (written to probs_out: softmax probabilities,
//         or the raw logits of forward() when `scale` == 0)
template <int F, int MODE>
__device__ __forceinline__ void process_chunk(LaneState<F>& st, const float* __restrict__ xs, const ChunkMeta& cm,
                                              uint32_t keepbits, float drop_scale, float scale, int lane,
                                              float* probs_out) {
    using D = Dim<F>;
    float aa[R], ab[R];
    float v[2 * R];
    // drop_scale: 1.25f = 1/(1-0.2) when dropout is active (ATen's noise.div_(1-p), train_lightning_ddp.py:60),
    // 1.0f when it is off (eval semantics / dropout_p == 0)
    const float DROP_SCALE = drop_scale;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        // keep the scheduler from hoisting every row's LDS.128 to the top (register pressure): rows are
        // software-pipelined two at a time
        if ((r & 1) == 0) asm volatile("" ::: "memory");
        float za = st.b1a, zb = st.b1b;
        const float4* x4 = reinterpret_cast<const float4*>(xs + r * D::FS);
#pragma unroll
        for (int q = 0; q < D::FS / 4; ++q) {
            const float4 xv = x4[q];
            const float xe[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int f = q * 4 + e;
                if (f < F) {
                    za = fmaf(xe[e], st.w1a[f], za);
                    zb = fmaf(xe[e], st.w1b[f], zb);
                }
            }
        }
        // relu with torch's NaN behaviour (threshold: x <= 0 ? 0 : x), then dropout scale
        float ha = (za <= 0.0f) ? 0.0f : za;
        float hb = (zb <= 0.0f) ? 0.0f : zb;
        if (MODE == 0) {
            ha *= ((keepbits >> (2 * r)) & 1u) ? DROP_SCALE : 0.0f;
            hb *= ((keepbits >> (2 * r + 1)) & 1u) ? DROP_SCALE : 0.0f;
        }
        aa[r] = ha;
        ab[r] = hb;
        v[2 * r + 0] = fmaf(hb, st.w2b0, ha * st.w2a0);
        v[2 * r + 1] = fmaf(hb, st.w2b1, ha * st.w2a1);
    }
    const float mine = reduce16_transpose(v, lane);
    const float other = __shfl_xor_sync(0xffffffffu, mine, 2);
    const int myrow = lane >> 2;
    const bool c1 = (lane >> 1) & 1;
    const float lg0 = (c1 ? other : mine) + st.b2_0;
    const float lg1 = (c1 ? mine : other) + st.b2_1;
    const int yv = __shfl_sync(0xffffffffu, cm.y, myrow);
    const int64_t rid = __shfl_sync(0xffffffffu, cm.rowid, myrow);
    const bool valid = rid >= 0;
    // log_softmax exactly as ATen orders it: (x - max) - log(sum exp(x - max))
    const float mx = fmaxf(lg0, lg1);
    const float s0 = lg0 - mx, s1 = lg1 - mx;
    const float lsum = logf(expf(s0) + expf(s1));
    const float lp0 = s0 - lsum, lp1 = s1 - lsum;
    float dl0 = 0.0f, dl1 = 0.0f;
    if (MODE == 2) {
        if (valid && (lane & 1) == 0)
            probs_out[rid * 2 + (c1 ? 1 : 0)] = (scale == 0.0f) ? (c1 ? lg1 : lg0) : expf(c1 ? lp1 : lp0);
        return;
    }
    const float lrow = -(yv ? lp1 : lp0);
    if (valid && (lane & 3) == 0) st.loss += lrow;
    if (MODE == 1) {
        // argmax returns the first maximal index; NaN counts as maximal (torch.argmax)
        const int pred = (lg1 > lg0 || (lg1 != lg1 && lg0 == lg0)) ? 1 : 0;
        if (valid && (lane & 3) == 0 && pred == yv) st.correct += 1;
        return;
    }
    if (valid) {
        dl0 = (expf(lp0) - (yv == 0 ? 1.0f : 0.0f)) * scale;
        dl1 = (expf(lp1) - (yv == 1 ? 1.0f : 0.0f)) * scale;
        if ((lane & 3) == 0) {
            st.gb2_0 += dl0;
            st.gb2_1 += dl1;
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if ((r & 1) == 0) asm volatile("" ::: "memory");
        const float d0 = __shfl_sync(0xffffffffu, dl0, 4 * r);
        const float d1 = __shfl_sync(0xffffffffu, dl1, 4 * r);
        st.gw2a0 = fmaf(d0, aa[r], st.gw2a0);
        st.gw2b0 = fmaf(d0, ab[r], st.gw2b0);
        st.gw2a1 = fmaf(d1, aa[r], st.gw2a1);
        st.gw2b1 = fmaf(d1, ab[r], st.gw2b1);
        // d a = dlogits . W2 ; dropout backward (x1.25 where kept) ; relu backward (where output > 0)
        const float ta = fmaf(d1, st.w2a1, d0 * st.w2a0);
        const float tb = fmaf(d1, st.w2b1, d0 * st.w2b0);
        const float dha = (aa[r] <= 0.0f) ? 0.0f : ta * DROP_SCALE;
        const float dhb = (ab[r] <= 0.0f) ? 0.0f : tb * DROP_SCALE;
        st.gb1a += dha;
        st.gb1b += dhb;
        const float4* x4 = reinterpret_cast<const float4*>(xs + r * D::FS);
#pragma unroll
        for (int q = 0; q < D::FS / 4; ++q) {
            const float4 xv = x4[q];
            const float xe[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int f = q * 4 + e;
                if (f < F) {
                    st.gw1a[f] = fmaf(dha, xe[e], st.gw1a[f]);
                    st.gw1b[f] = fmaf(dhb, xe[e], st.gw1b[f]);
                }
            }
        }
    }
}

// Adam for one element, torch.optim.Adam single-tensor semantics (train_lightning_ddp.py:88).
struct AdamScalars {
    float w_lerp;       // float(1 - beta1)
    float beta2;        // float(beta2)
    float one_m_beta2;  // float(1 - beta2)
    float neg_step;     // float(-(lr / bias_correction1))
    float bc2_sqrt;     // float(sqrt(bias_correction2))
    float eps;
    int sgd;            // WX_OPT_SGD: p += neg_step * g with neg_step = -lr, moments untouched
};

__device__ __forceinline__ AdamScalars adam_scalars(const wx_model_desc& d, int64_t step /*1-based*/) {
    // torch computes the bias corrections and step size in Python doubles and hands fp32 scalars to
    // the element-wise kernels; same here.
    AdamScalars a;
    a.sgd = d.optimizer == WX_OPT_SGD;
    if (a.sgd) {
        a.w_lerp = a.beta2 = a.one_m_beta2 = a.bc2_sqrt = a.eps = 0.0f;
        a.neg_step = float(-d.lr);
        return a;
    }
    const double b1 = d.beta1, b2 = d.beta2;
    const double bc1 = 1.0 - pow(b1, double(step));
    const double bc2 = 1.0 - pow(b2, double(step));
    a.w_lerp = float(1.0 - b1);
    a.beta2 = float(b2);
    a.one_m_beta2 = float(1.0 - b2);
    a.neg_step = float(-(d.lr / bc1));
    a.bc2_sqrt = float(sqrt(bc2));
    a.eps = float(d.eps);
    return a;
}

__device__ __forceinline__ void adam_apply(float& p, float g, float& m, float& v, const AdamScalars& a) {
    if (a.sgd) {
        p = fmaf(a.neg_step, g, p);                            // param.add_(grad, alpha=-lr), one rounding like ATen's fused add
        return;
    }
    m = fmaf(a.w_lerp, g - m, m);                              // exp_avg.lerp_(grad, 1-beta1)
    v = v * a.beta2;                                           // exp_avg_sq.mul_(beta2)
    v = v + (a.one_m_beta2 * g) * g;                           //   .addcmul_(grad, grad, value=1-beta2)
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;         // (exp_avg_sq.sqrt() / bc2_sqrt).add_(eps)
    p = p + (a.neg_step * m) / denom;                          // param.addcdiv_(exp_avg, denom, value=-step_size)
}

}  // namespace wxk
