// Kernel templates: K2 (fused forward + cross-entropy + backward), K3 (Adam), K5 (eval / predict).
// Replaces /root/reference/jobs/train_lightning_ddp.py:57-88 (model, training_step, validation_step,
// configure_optimizers) and, in the raw-column loader, /root/reference/jobs/preprocess.py:38-41.
#pragma once
#include <math.h>
#include <string.h>

#include "mlp_step.cuh"

namespace wxk {

constexpr int THREADS = 256;
constexpr int WARPS = THREADS / 32;

enum Mode { MODE_TRAIN = 0, MODE_EVAL = 1, MODE_PREDICT = 2 };
enum Tail { TAIL_GRADS = 0, TAIL_ADAM = 1 };

struct GridArgs {
    // model
    const float* params;   // [P] read at kernel start
    wx_model_desc desc;
    // data
    int loader;
    const float* x;
    const int64_t* y;
    const int64_t* row_idx;
    int64_t row0;
    int64_t batch;
    WxCols cols;
    const double* stats;
    int impute;
    // dropout
    int dropout_mode;
    const uint64_t* keep_mask;
    uint64_t seed;
    uint64_t step;          // Philox step word when state == nullptr
    float scale;
    // reduction scratch
    float* part;            // [grid][SLAB] fp32 (train) or [grid][2] doubles (eval)
    unsigned int* ticket;
    // outputs
    int tail;
    float* out;             // TAIL_GRADS: [P+1]
    float* p_rw;            // TAIL_ADAM
    float* m;
    float* v;
    int64_t* state;
    float* loss_out;
    double* eval_out;       // MODE_EVAL: [2]
    long long* dbg;         // optional per-CTA timestamps (wxdev_cta_times)
    float* probs_out;       // MODE_PREDICT
    wxpeer::PeerArgs peer;  // world > 1: cross-rank sum inside the tail (TAIL_ADAM only)
};

// smem bytes per warp for the staging buffers of a loader
template <int F>
__host__ __device__ constexpr int stage_bytes_per_warp(int loader) {
    return loader == LOAD_RAW_F64 ? (2 * R * F * 8 + R * Dim<F>::FS * 4) : (2 * R * Dim<F>::FS * 4);
}
template <int F>
__host__ __device__ constexpr int smem_bytes(int loader, int mode) {
    const int stage = WARPS * stage_bytes_per_warp<F>(loader);
    const int slab = mode == MODE_TRAIN ? Dim<F>::SLAB * 4 : 256;
    return (stage > slab ? stage : slab) + 2 * WX_MAX_FEATURES * 8 /* mean/std */ + 64;
}

// ---- staging ------------------------------------------------------------------------------------
template <int F>
__device__ __forceinline__ ChunkMeta load_meta(const GridArgs& a, int64_t chunk, int lane) {
    ChunkMeta cm;
    cm.rowid = -1;
    cm.y = 0;
    cm.mlo = cm.mhi = 0u;
    const int64_t pos = chunk * R + lane;
    if (lane < R && pos < a.batch) {
        cm.rowid = a.row_idx ? a.row_idx[pos] : a.row0 + pos;
        if (a.y) cm.y = int(a.y[cm.rowid]);
        if (a.dropout_mode == WX_DROPOUT_MASK && a.keep_mask) {
            const uint64_t mk = a.keep_mask[pos];
            cm.mlo = uint32_t(mk);
            cm.mhi = uint32_t(mk >> 32);
        }
    }
    return cm;
}

template <int F, int LOADER>
__device__ __forceinline__ void stage_chunk(const GridArgs& a, int64_t chunk, const ChunkMeta& cm, void* stg, int lane) {
    using D = Dim<F>;
    const int64_t pos0 = chunk * R;
    if (LOADER == LOAD_ROWMAJOR) {
        float* dst = static_cast<float*>(stg);
        constexpr int ITERS = (R * F + 31) / 32;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int e = it * 32 + lane;
            const bool act = e < R * F;
            const int r = act ? e / F : 0;
            const int f = e - r * F;
            int64_t rid;
            if (a.row_idx) {
                rid = __shfl_sync(0xffffffffu, cm.rowid, r);
            } else {
                rid = (pos0 + r < a.batch) ? a.row0 + pos0 + r : -1;
            }
            if (act) {
                if (rid >= 0)
                    cp_async4(dst + r * D::FS + f, a.x + rid * F + f);
                else
                    dst[r * D::FS + f] = 0.0f;
            }
        }
    } else {
        // raw columns: element e -> (column j = e / R, chunk row r = e % R); 8 consecutive rows of one
        // column are one 32 B (fp32) / 64 B (fp64) segment
        constexpr int ITERS = (R * F + 31) / 32;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int e = it * 32 + lane;
            if (e < R * F) {
                const int j = e / R, r = e % R;
                const bool in = pos0 + r < a.batch;
                const int64_t row = a.row0 + pos0 + r;
                if (LOADER == LOAD_RAW_F32) {
                    float* dst = static_cast<float*>(stg) + r * D::FS + j;
                    if (in)
                        cp_async4(dst, static_cast<const float*>(a.cols.col[j]) + row);
                    else
                        *dst = 0.0f;
                } else {
                    double* dst = static_cast<double*>(stg) + r * F + j;
                    if (in)
                        cp_async8(dst, static_cast<const double*>(a.cols.col[j]) + row);
                    else
                        *dst = 0.0;
                }
            }
        }
    }
}

// standardise a staged raw chunk (preprocess.py:38-41 + the fp32 cast at train_lightning_ddp.py:45)
template <int F, int LOADER>
__device__ __forceinline__ void normalize_chunk(const GridArgs& a, const void* stg, float* xs, const double* s_mean,
                                                const double* s_std, int64_t chunk, int lane) {
    using D = Dim<F>;
    constexpr int ITERS = (R * F + 31) / 32;
    const int64_t pos0 = chunk * R;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int e = it * 32 + lane;
        if (e < R * F) {
            const int j = e / R, r = e % R;
            double x;
            if (LOADER == LOAD_RAW_F32)
                x = double(static_cast<const float*>(stg)[r * D::FS + j]);
            else
                x = static_cast<const double*>(stg)[r * F + j];
            float v = __double2float_rn((x - s_mean[j]) / s_std[j]);
            if (isnan(x)) v = (a.impute == WX_IMPUTE_MEAN) ? 0.0f : __int_as_float(0x7fc00000);
            if (pos0 + r >= a.batch) v = 0.0f;
            xs[r * D::FS + j] = v;
        }
    }
}

// ---- the grid-wide step kernel -------------------------------------------------------------------
template <int F, int MODE, int LOADER>
__global__ void __launch_bounds__(THREADS) mlp_grid_kernel(const GridArgs a) {
    using D = Dim<F>;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    constexpr int stage_per_warp = stage_bytes_per_warp<F>(LOADER);
    constexpr int stage_total = WARPS * stage_per_warp;
    constexpr int slab_bytes = MODE == MODE_TRAIN ? D::SLAB * 4 : 256;
    constexpr int region = stage_total > slab_bytes ? stage_total : slab_bytes;
    double* s_mean = reinterpret_cast<double*>(smem + region);
    double* s_std = s_mean + WX_MAX_FEATURES;
    unsigned char* my = smem + warp * stage_per_warp;
    // per-warp views: two staging buffers `stg_stride` bytes apart (+ one fp32 chunk for the f64 loader)
    constexpr int stg_stride = (LOADER == LOAD_RAW_F64) ? R * F * 8 : R * D::FS * 4;
    float* xs_single = reinterpret_cast<float*>(my + 2 * R * F * 8);   // LOAD_RAW_F64 only
    if (LOADER != LOAD_ROWMAJOR) {
        if (tid < F) {
            s_mean[tid] = a.stats[tid * 4 + 1];
            s_std[tid] = a.stats[tid * 4 + 3];
        }
        __syncthreads();
    }

    if (a.dbg != nullptr && tid == 0) {
        unsigned smid;
        long long t;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.dbg[blockIdx.x * 4 + 0] = smid;
        a.dbg[blockIdx.x * 4 + 1] = t;
    }
    LaneState<F> st;
    st.load_weights(a.params, lane);
    st.zero_grads();

    const uint64_t step_word = (a.state != nullptr) ? uint64_t(a.state[0]) : a.step;
    const float drop_scale = (a.dropout_mode == WX_DROPOUT_OFF) ? 1.0f : 1.25f;
    const uint2 key = make_uint2(uint32_t(a.seed), uint32_t(a.seed >> 32));

    const int64_t nchunks = (a.batch + R - 1) / R;
    const int64_t gwarps = int64_t(gridDim.x) * WARPS;
    int64_t c = int64_t(blockIdx.x) * WARPS + warp;
    ChunkMeta nxt = load_meta<F>(a, c < nchunks ? c : 0, c < nchunks ? lane : 32);
    if (c < nchunks) stage_chunk<F, LOADER>(a, c, nxt, my, lane);
    cp_async_commit();
    int buf = 0;
    for (; c < nchunks; c += gwarps) {
        const ChunkMeta cur = nxt;
        const int64_t cn = c + gwarps;
        if (cn < nchunks) {
            nxt = load_meta<F>(a, cn, lane);
            stage_chunk<F, LOADER>(a, cn, nxt, my + (buf ^ 1) * stg_stride, lane);
        }
        cp_async_commit();
        cp_async_wait<1>();
        __syncwarp();
        unsigned char* cur_stg = my + buf * stg_stride;
        const float* xs = reinterpret_cast<const float*>(cur_stg);
        if (LOADER == LOAD_RAW_F32) {
            normalize_chunk<F, LOADER>(a, cur_stg, reinterpret_cast<float*>(cur_stg), s_mean, s_std, c, lane);
            __syncwarp();
        } else if (LOADER == LOAD_RAW_F64) {
            normalize_chunk<F, LOADER>(a, cur_stg, xs_single, s_mean, s_std, c, lane);
            __syncwarp();
            xs = xs_single;
        }
        uint32_t keepbits = 0xffffu;
        if (MODE == MODE_TRAIN)
            keepbits = chunk_keep_bits(a.dropout_mode, cur, lane, c * R, key, uint32_t(step_word), uint32_t(step_word >> 32));
        process_chunk<F, MODE>(st, xs, cur, keepbits, drop_scale, a.scale, lane, a.probs_out);
        __syncwarp();
        buf ^= 1;
    }
    cp_async_wait<0>();
    if (MODE == MODE_PREDICT) return;

    __shared__ bool s_last;
    if (a.dbg != nullptr) {
        __syncthreads();
        if (tid == 0) {
            long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            a.dbg[blockIdx.x * 4 + 2] = t;
        }
    }
    __shared__ AdamScalars s_adam;
    __syncthreads();   // staging memory is dead from here on; the slab aliases it

    if (MODE == MODE_EVAL) {
        double* red = reinterpret_cast<double*>(smem);   // [WARPS][2]
        const double ls = warp_sum(double(st.loss));
        const double cs = warp_sum(double(st.correct));
        if (lane == 0) {
            red[warp * 2 + 0] = ls;
            red[warp * 2 + 1] = cs;
        }
        __syncthreads();
        double* part = reinterpret_cast<double*>(a.part);
        if (tid == 0) {
            double l = 0.0, k = 0.0;
            for (int w = 0; w < WARPS; ++w) {
                l += red[w * 2];
                k += red[w * 2 + 1];
            }
            part[blockIdx.x * 2 + 0] = l;
            part[blockIdx.x * 2 + 1] = k;
            __threadfence();
            s_last = atomicAdd(a.ticket, 1u) == gridDim.x - 1;
        }
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        if (tid == 0) {
            double l = 0.0, k = 0.0;
            for (unsigned b = 0; b < gridDim.x; ++b) {
                l += __ldcg(part + b * 2);
                k += __ldcg(part + b * 2 + 1);
            }
            a.eval_out[0] = l;
            a.eval_out[1] = k;
            *a.ticket = 0u;
        }
        return;
    }

    // MODE_TRAIN: warps fold their register sums into the CTA slab in warp order (deterministic)
    float* slab = reinterpret_cast<float*>(smem);
    for (int w = 0; w < WARPS; ++w) {
        if (warp == w) st.flush_to_slab(slab, lane, w == 0);
        __syncthreads();
    }
    float* mypart = a.part + size_t(blockIdx.x) * D::SLAB;
    for (int i = tid; i < D::P + 1; i += THREADS) mypart[i] = slab[i];
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(a.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (a.dbg != nullptr && tid == 0) {
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.dbg[blockIdx.x * 4 + 3] = t;
    }
    if (!s_last) return;
    __threadfence();

    // last CTA: sum the slabs in CTA order, then either publish grads or apply Adam in place
    if (a.tail == TAIL_ADAM && tid == 0) s_adam = adam_scalars(a.desc, a.state[0] + 1);
    __syncthreads();
    // Each thread owns up to GPT groups of 4 consecutive outputs and walks the slabs in CTA order with
    // independent 128-bit loads in flight (the order per element is b = 0,1,2,... -> deterministic).
    const unsigned nparts = gridDim.x;
    constexpr int NG = D::SLAB / 4;
    constexpr int GPT = (NG + THREADS - 1) / THREADS;
    float4 acc[GPT];
#pragma unroll
    for (int k = 0; k < GPT; ++k) acc[k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const float4* q4 = reinterpret_cast<const float4*>(a.part);
#pragma unroll 8
    for (unsigned b = 0; b < nparts; ++b) {
#pragma unroll
        for (int k = 0; k < GPT; ++k) {
            int g = tid + k * THREADS;
            g = g < NG ? g : NG - 1;          // surplus threads re-read the last group (branch-free, result unused)
            const float4 t = __ldcg(q4 + size_t(b) * NG + g);
            acc[k].x += t.x;
            acc[k].y += t.y;
            acc[k].z += t.z;
            acc[k].w += t.w;
        }
    }
#pragma unroll
    for (int k = 0; k < GPT; ++k) {
        const float vals[4] = {acc[k].x, acc[k].y, acc[k].z, acc[k].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int i = 4 * (tid + k * THREADS) + e;
            if (i > D::P) continue;
            float s = vals[e];
            if (i == D::P) s *= a.scale;   // loss: rows were summed unscaled
            if (a.tail == TAIL_GRADS) {
                a.out[i] = s;
            } else if (i < D::P) {
                float p = a.p_rw[i], m = a.m[i], v = a.v[i];
                adam_apply(p, s, m, v, s_adam);
                a.p_rw[i] = p;
                a.m[i] = m;
                a.v[i] = v;
            } else {
                *a.loss_out = s;
            }
        }
    }
    if (tid == 0) {
        if (a.dbg != nullptr) {
            long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            a.dbg[blockIdx.x * 4 + 3] = t;
        }
        *a.ticket = 0u;
        if (a.tail == TAIL_ADAM) a.state[0] += 1;
    }
}

// ---- stand-alone Adam (K3) -----------------------------------------------------------------------
static __global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                   AdamScalars a, int vec_ok) {
    const int64_t stride = int64_t(gridDim.x) * blockDim.x;
    const int64_t t0 = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (vec_ok) {
        const int64_t n4 = n >> 2;
        float4* p4 = reinterpret_cast<float4*>(p);
        const float4* g4 = reinterpret_cast<const float4*>(g);
        float4* m4 = reinterpret_cast<float4*>(m);
        float4* v4 = reinterpret_cast<float4*>(v);
        for (int64_t i = t0; i < n4; i += stride) {
            float4 pp = p4[i], gg = __ldg(g4 + i), mm = m4[i], vv = v4[i];
            adam_apply(pp.x, gg.x, mm.x, vv.x, a);
            adam_apply(pp.y, gg.y, mm.y, vv.y, a);
            adam_apply(pp.z, gg.z, mm.z, vv.z, a);
            adam_apply(pp.w, gg.w, mm.w, vv.w, a);
            p4[i] = pp;
            m4[i] = mm;
            v4[i] = vv;
        }
        for (int64_t i = (n4 << 2) + t0; i < n; i += stride) {
            float pp = p[i], mm = m[i], vv = v[i];
            adam_apply(pp, g[i], mm, vv, a);
            p[i] = pp;
            m[i] = mm;
            v[i] = vv;
        }
    } else {
        for (int64_t i = t0; i < n; i += stride) {
            float pp = p[i], mm = m[i], vv = v[i];
            adam_apply(pp, g[i], mm, vv, a);
            p[i] = pp;
            m[i] = mm;
            v[i] = vv;
        }
    }
}

// ---- many tiny steps in one launch (the reference's batch-4 regime) -------------------------------
// One CTA; params / m / v live in shared memory across steps; each warp owns at most one chunk of a
// step (batch <= WARPS * R); next step's rows are prefetched while the current one is computed.
struct SmallArgs {
    float* params;
    float* m;
    float* v;
    int64_t* state;
    wx_model_desc desc;
    const float* x;
    const int64_t* y;
    const int64_t* row_idx;
    int64_t row0;
    int64_t batch;
    int64_t n_steps;
    int dropout_mode;
    const uint64_t* keep_mask;
    uint64_t seed;
    float* loss_out;
    // resident multi-CTA variant only:
    float* part;                    // [2][gridDim.x][slab_stride] CTA partial sums, double-buffered by step parity
    unsigned long long* bar;        // grid barrier counter, zero at launch
    int* err;                       // sticky status word
    wxpeer::PeerArgs peer;          // world > 1: the [P+1] sums are exchanged over NVLink inside every step (one CTA only)
};

template <int F>
__global__ void __launch_bounds__(THREADS, 1) mlp_small_steps_kernel(const SmallArgs sa) {
    using D = Dim<F>;
    extern __shared__ __align__(16) unsigned char smem[];
    float* s_p = reinterpret_cast<float*>(smem);
    float* s_m = s_p + D::SLAB;
    float* s_v = s_m + D::SLAB;
    float* slab = s_v + D::SLAB;
    float* stage_base = slab + D::SLAB;                    // [WARPS][2][R*FS]
    __shared__ AdamScalars s_adam;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < D::P; i += THREADS) {
        s_p[i] = sa.params[i];
        s_m[i] = sa.m[i];
        s_v[i] = sa.v[i];
    }
    __syncthreads();
    const int64_t step0 = sa.state[0];
    const uint2 key = make_uint2(uint32_t(sa.seed), uint32_t(sa.seed >> 32));
    const int nchunks = int((sa.batch + R - 1) / R);       // <= WARPS
    const bool has_chunk = warp < nchunks;
    const bool ddp = sa.peer.world > 1;                    // mean over world * batch rows, as DDP's averaged gradients give
    const float scale = 1.0f / (float(sa.batch) * float(ddp ? sa.peer.world : 1));
    int64_t done = sa.n_steps;                             // steps actually taken (fewer if a peer stops arriving)
    float* stg0 = stage_base + warp * 2 * R * D::FS;

    // a GridArgs view so load_meta / stage_chunk can be reused; row window moves with the step
    GridArgs ga;
    ga.loader = LOAD_ROWMAJOR;
    ga.x = sa.x;
    ga.y = sa.y;
    ga.batch = sa.batch;
    ga.dropout_mode = sa.dropout_mode;

    auto window = [&](int64_t s) {
        ga.row_idx = sa.row_idx ? sa.row_idx + s * sa.batch : nullptr;
        ga.row0 = sa.row0 + s * sa.batch;
        ga.keep_mask = sa.keep_mask ? sa.keep_mask + s * sa.batch : nullptr;
    };
    ChunkMeta nxt;
    nxt.rowid = -1;
    nxt.y = 0;
    nxt.mlo = nxt.mhi = 0;
    if (has_chunk && sa.n_steps > 0) {
        window(0);
        nxt = load_meta<F>(ga, warp, lane);
        stage_chunk<F, LOAD_ROWMAJOR>(ga, warp, nxt, stg0, lane);
    }
    cp_async_commit();
    int buf = 0;
    for (int64_t s = 0; s < sa.n_steps; ++s) {
        const ChunkMeta cur = nxt;
        if (has_chunk && s + 1 < sa.n_steps) {
            window(s + 1);
            nxt = load_meta<F>(ga, warp, lane);
            stage_chunk<F, LOAD_ROWMAJOR>(ga, warp, nxt, stg0 + (buf ^ 1) * R * D::FS, lane);
        }
        cp_async_commit();
        if (has_chunk) {
            LaneState<F> st;
            st.load_weights(s_p, lane);
            st.zero_grads();
            cp_async_wait<1>();
            __syncwarp();
            const uint64_t step_word = uint64_t(step0 + s);
            const uint32_t keepbits = chunk_keep_bits(sa.dropout_mode, cur, lane, int64_t(warp) * R, key,
                                                      uint32_t(step_word), uint32_t(step_word >> 32));
            process_chunk<F, MODE_TRAIN>(st, stg0 + buf * R * D::FS, cur, keepbits,
                                         (sa.dropout_mode == WX_DROPOUT_OFF) ? 1.0f : 1.25f, scale, lane, nullptr);
            // fold into the slab in warp order
            for (int w = 0; w < nchunks; ++w) {
                if (warp == w) st.flush_to_slab(slab, lane, w == 0);
                asm volatile("bar.sync 1, %0;" ::"r"(nchunks * 32) : "memory");
            }
        }
        if (tid == 0) s_adam = adam_scalars(sa.desc, step0 + s + 1);
        __syncthreads();
        if (ddp) {
            // DDP's bucket all-reduce + the sync_dist loss (train_lightning_ddp.py:70,:136) without leaving the kernel:
            // the [P+1] sums cross NVLink once per step (csrc/peer.cuh).  A peer that stops arriving ends the launch here
            // with the step NOT applied; the status word tells the host.
            if (!wxpeer::allreduce_cta(slab, D::P + 1, sa.peer, uint64_t(step0 + s + 1), tid, THREADS)) {
                if (tid == 0 && sa.err != nullptr) atomicExch(sa.err, 20);
                done = s;
                break;
            }
        }
        for (int i = tid; i < D::P; i += THREADS) adam_apply(s_p[i], slab[i], s_m[i], s_v[i], s_adam);
        if (tid == 0) sa.loss_out[s] = slab[D::P] * scale;
        __syncthreads();
        buf ^= 1;
    }
    cp_async_wait<0>();
    for (int i = tid; i < D::P; i += THREADS) {
        sa.params[i] = s_p[i];
        sa.m[i] = s_m[i];
        sa.v[i] = s_v[i];
    }
    if (tid == 0) sa.state[0] = step0 + done;
}


// The same many-steps-per-launch idea for batches that need more than one CTA (64 < batch <= 8192 or so: the
// latency-bound end of the batch sweep, BASELINE.json configs[4]).  gridDim.x CTAs stay resident (cooperative launch) and
// every one of them keeps params / m / v in shared memory.  Per step: each CTA processes its contiguous range of 8-row
// chunks (warps double-buffer their chunks with cp.async), folds its warps into a CTA slab in warp order, publishes the
// slab to global memory (two buffers, by step parity), crosses ONE grid barrier, then every CTA adds all slabs in CTA
// order and applies Adam to its own copy -- identical arithmetic everywhere, so the copies never diverge and no second
// barrier or broadcast is needed.  CTA 0 writes the losses and, at the end, the parameters.
template <int F>
__global__ void __launch_bounds__(THREADS, 1) mlp_resident_steps_kernel(const SmallArgs sa) {
    using D = Dim<F>;
    extern __shared__ __align__(16) unsigned char smem[];
    float* s_p = reinterpret_cast<float*>(smem);
    float* s_m = s_p + D::SLAB;
    float* s_v = s_m + D::SLAB;
    float* slab = s_v + D::SLAB;
    float* stage_base = slab + D::SLAB;                    // [WARPS][2][R*FS]
    __shared__ AdamScalars s_adam;
    __shared__ int s_fail;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int G = int(gridDim.x);
    for (int i = tid; i < D::P; i += THREADS) {
        s_p[i] = sa.params[i];
        s_m[i] = sa.m[i];
        s_v[i] = sa.v[i];
    }
    if (tid == 0) s_fail = 0;
    __syncthreads();
    const int64_t step0 = sa.state[0];
    const uint2 key = make_uint2(uint32_t(sa.seed), uint32_t(sa.seed >> 32));
    const int64_t nchunks = (sa.batch + R - 1) / R;
    const int64_t cpc = (nchunks + G - 1) / G;             // chunks per CTA (the host sizes G so that every CTA has work)
    const int64_t c_lo = int64_t(blockIdx.x) * cpc;
    const int64_t c_hi = (c_lo + cpc < nchunks) ? c_lo + cpc : nchunks;
    const int n_active = int((c_hi - c_lo) < WARPS ? (c_hi - c_lo > 0 ? c_hi - c_lo : 0) : WARPS);
    const float scale = 1.0f / float(sa.batch);
    const float ds = (sa.dropout_mode == WX_DROPOUT_OFF) ? 1.0f : 1.25f;
    float* stg0 = stage_base + warp * 2 * R * D::FS;
    constexpr int SLAB_STRIDE = (D::P + 1 + 3) & ~3;

    GridArgs ga;
    ga.loader = LOAD_ROWMAJOR;
    ga.x = sa.x;
    ga.y = sa.y;
    ga.batch = sa.batch;
    ga.dropout_mode = sa.dropout_mode;

    for (int64_t s = 0; s < sa.n_steps; ++s) {
        ga.row_idx = sa.row_idx ? sa.row_idx + s * sa.batch : nullptr;
        ga.row0 = sa.row0 + s * sa.batch;
        ga.keep_mask = sa.keep_mask ? sa.keep_mask + s * sa.batch : nullptr;
        if (warp < n_active) {
            LaneState<F> st;
            st.load_weights(s_p, lane);
            st.zero_grads();
            int64_t c = c_lo + warp;
            ChunkMeta nxt = load_meta<F>(ga, c, lane);
            stage_chunk<F, LOAD_ROWMAJOR>(ga, c, nxt, stg0, lane);
            cp_async_commit();
            int buf = 0;
            const uint64_t step_word = uint64_t(step0 + s);
            for (; c < c_hi; c += WARPS) {
                const ChunkMeta cur = nxt;
                const int64_t cn = c + WARPS;
                if (cn < c_hi) {
                    nxt = load_meta<F>(ga, cn, lane);
                    stage_chunk<F, LOAD_ROWMAJOR>(ga, cn, nxt, stg0 + (buf ^ 1) * R * D::FS, lane);
                }
                cp_async_commit();
                cp_async_wait<1>();
                __syncwarp();
                const uint32_t keepbits = chunk_keep_bits(sa.dropout_mode, cur, lane, c * R, key, uint32_t(step_word),
                                                          uint32_t(step_word >> 32));
                process_chunk<F, MODE_TRAIN>(st, stg0 + buf * R * D::FS, cur, keepbits, ds, scale, lane, nullptr);
                __syncwarp();
                buf ^= 1;
            }
            cp_async_wait<0>();
            for (int w = 0; w < n_active; ++w) {                      // fold into the CTA slab in warp order
                if (warp == w) st.flush_to_slab(slab, lane, w == 0);
                asm volatile("bar.sync 1, %0;" ::"r"(n_active * 32) : "memory");
            }
        }
        __syncthreads();
        if (G > 1) {
            float* mine = sa.part + (size_t(s & 1) * G + blockIdx.x) * SLAB_STRIDE;
            for (int i = tid; i < D::P + 1; i += THREADS) mine[i] = slab[i];
            __threadfence();
            __syncthreads();
            if (tid == 0) {
                atomicAdd(sa.bar, 1ull);
                const unsigned long long target = (unsigned long long)(s + 1) * G;
                bool ok = false;
                for (uint32_t spin = 0; spin < (1u << 26); ++spin) {
                    unsigned long long seen;
                    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(seen) : "l"(sa.bar) : "memory");
                    if (seen >= target) { ok = true; break; }
                }
                if (!ok) {
                    s_fail = 1;
                    atomicExch(sa.err, 30);                          // grid barrier never filled: CTAs not co-resident?
                }
            }
            __syncthreads();
            if (s_fail) return;
            const float* all = sa.part + size_t(s & 1) * G * SLAB_STRIDE;
            for (int i = tid; i < D::P + 1; i += THREADS) {
                float t = 0.0f;
                for (int g = 0; g < G; ++g) t += __ldcg(all + size_t(g) * SLAB_STRIDE + i);   // CTA order: same sum everywhere
                slab[i] = t;
            }
        }
        if (tid == 0) s_adam = adam_scalars(sa.desc, step0 + s + 1);
        __syncthreads();
        for (int i = tid; i < D::P; i += THREADS) adam_apply(s_p[i], slab[i], s_m[i], s_v[i], s_adam);
        if (tid == 0 && blockIdx.x == 0) sa.loss_out[s] = slab[D::P] * scale;
        __syncthreads();
    }
    if (blockIdx.x == 0) {
        for (int i = tid; i < D::P; i += THREADS) {
            sa.params[i] = s_p[i];
            sa.m[i] = s_m[i];
            sa.v[i] = s_v[i];
        }
        if (tid == 0) sa.state[0] = step0 + sa.n_steps;
    }
}


typedef void (*grid_kernel_t)(const GridArgs);
typedef void (*small_kernel_t)(const SmallArgs);

// Instantiations are spread over four translation units (mlp_inst.cu built with -DWX_F_GROUP=0..3,
// feature counts 8g+1 .. 8g+8) so the library builds in parallel.
grid_kernel_t pick_grid_kernel_g0(int f, int mode, int loader, int* smem);
grid_kernel_t pick_grid_kernel_g1(int f, int mode, int loader, int* smem);
grid_kernel_t pick_grid_kernel_g2(int f, int mode, int loader, int* smem);
grid_kernel_t pick_grid_kernel_g3(int f, int mode, int loader, int* smem);
small_kernel_t pick_small_kernel_g0(int f, int* smem);
small_kernel_t pick_small_kernel_g1(int f, int* smem);
small_kernel_t pick_small_kernel_g2(int f, int* smem);
small_kernel_t pick_small_kernel_g3(int f, int* smem);
small_kernel_t pick_resident_kernel_g0(int f, int* smem);
small_kernel_t pick_resident_kernel_g1(int f, int* smem);
small_kernel_t pick_resident_kernel_g2(int f, int* smem);
small_kernel_t pick_resident_kernel_g3(int f, int* smem);

}  // namespace wxk
