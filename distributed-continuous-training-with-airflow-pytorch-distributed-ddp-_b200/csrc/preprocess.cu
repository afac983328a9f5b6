// K1: label encode, column statistics, standardise + pack, imputation index list.
// Replaces the arithmetic of /root/reference/jobs/preprocess.py:23-41 and the casts at
// /root/reference/jobs/train_lightning_ddp.py:45-46.  All HBM-bound streaming kernels.
#include "common.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// label encode (preprocess.py:23-25)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) label_encode_kernel(const int32_t* __restrict__ off,
                                                          const uint8_t* __restrict__ data,
                                                          const uint8_t* __restrict__ valid, int64_t n,
                                                          int64_t* __restrict__ out64, int32_t* __restrict__ out32) {
    const int64_t stride = int64_t(gridDim.x) * blockDim.x;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int32_t a = off[i], b = off[i + 1];
        int32_t lab = 0;
        if (b - a == 4) {
            bool ok = valid == nullptr || ((valid[i >> 3] >> (i & 7)) & 1);
            if (ok) {
                const uint8_t* s = data + a;
                lab = (s[0] == 'r' && s[1] == 'a' && s[2] == 'i' && s[3] == 'n') ? 1 : 0;
            }
        }
        out64[i] = lab;
        if (out32) out32[i] = lab;
    }
}

// ------------------------------------------------------------------------------------------
// column statistics (preprocess.py:32-36): per-thread two-pass groups -> Chan merges, all float64
// ------------------------------------------------------------------------------------------
struct Moments {
    double n, mean, m2;
};

__device__ __forceinline__ Moments chan_merge(const Moments& a, const Moments& b) {
    if (b.n == 0.0) return a;
    if (a.n == 0.0) return b;
    Moments r;
    r.n = a.n + b.n;
    const double d = b.mean - a.mean;
    r.mean = a.mean + d * (b.n / r.n);
    r.m2 = a.m2 + b.m2 + d * d * (a.n * b.n / r.n);
    return r;
}

__device__ __forceinline__ Moments shfl_xor_m(const Moments& a, int o) {
    Moments r;
    r.n = __shfl_xor_sync(0xffffffffu, a.n, o);
    r.mean = __shfl_xor_sync(0xffffffffu, a.mean, o);
    r.m2 = __shfl_xor_sync(0xffffffffu, a.m2, o);
    return r;
}

__device__ __forceinline__ double finalize_std(const Moments& m) {
    if (m.n < 2.0) return __longlong_as_double(0x7ff8000000000000LL);   // Spark: stddev_samp of <2 rows is null
    const double sd = sqrt(m.m2 / (m.n - 1.0));
    return (sd != 0.0) ? sd : 1.0;                                       // preprocess.py:36
}

constexpr int STATS_THREADS = 256;
constexpr int STATS_G = 8;   // independent loads per thread per iteration

template <typename T>
__global__ void __launch_bounds__(STATS_THREADS) col_stats_kernel(WxCols cols, int n_feat, int64_t n,
                                                                  double* __restrict__ parts,
                                                                  unsigned int* __restrict__ ticket,
                                                                  double* __restrict__ stats_out) {
    const int j = blockIdx.y;
    const T* __restrict__ col = static_cast<const T*>(cols.col[j]);
    const uint8_t* __restrict__ vb = cols.valid[j];
    const int tid = threadIdx.x;
    Moments run = {0.0, 0.0, 0.0};
    const int64_t tile = int64_t(STATS_THREADS) * STATS_G;
    for (int64_t base = int64_t(blockIdx.x) * tile; base < n; base += int64_t(gridDim.x) * tile) {
        double x[STATS_G];
        bool ok[STATS_G];
#pragma unroll
        for (int k = 0; k < STATS_G; ++k) {
            const int64_t i = base + tid + int64_t(k) * STATS_THREADS;
            ok[k] = i < n;
            x[k] = ok[k] ? double(col[i]) : 0.0;
            if (ok[k] && vb != nullptr) ok[k] = (vb[i >> 3] >> (i & 7)) & 1;
            ok[k] = ok[k] && !isnan(x[k]);
        }
        double c = 0.0, s = 0.0;
#pragma unroll
        for (int k = 0; k < STATS_G; ++k)
            if (ok[k]) {
                c += 1.0;
                s += x[k];
            }
        if (c > 0.0) {
            Moments g;
            g.n = c;
            g.mean = s / c;
            g.m2 = 0.0;
#pragma unroll
            for (int k = 0; k < STATS_G; ++k)
                if (ok[k]) {
                    const double d = x[k] - g.mean;
                    g.m2 += d * d;
                }
            run = chan_merge(run, g);
        }
    }
    // fixed-shape tree inside the warp, then warps in order: deterministic for a given grid
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        Moments other = shfl_xor_m(run, o);
        // keep operand order independent of the lane so both partners compute the same value
        run = (tid & o) ? chan_merge(other, run) : chan_merge(run, other);
    }
    __shared__ Moments wm[STATS_THREADS / 32];
    __shared__ bool is_last;
    if ((tid & 31) == 0) wm[tid >> 5] = run;
    __syncthreads();
    if (tid == 0) {
        Moments b = wm[0];
        for (int w = 1; w < STATS_THREADS / 32; ++w) b = chan_merge(b, wm[w]);
        double* p = parts + (size_t(j) * gridDim.x + blockIdx.x) * 3;
        p[0] = b.n;
        p[1] = b.mean;
        p[2] = b.m2;
        __threadfence();
        const unsigned int total = gridDim.x * gridDim.y;
        is_last = (atomicAdd(ticket, 1u) == total - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    if (tid < n_feat) {
        const volatile double* p = parts + size_t(tid) * gridDim.x * 3;
        Moments acc = {0.0, 0.0, 0.0};
        for (unsigned int b = 0; b < gridDim.x; ++b) {
            Moments q = {p[b * 3 + 0], p[b * 3 + 1], p[b * 3 + 2]};
            acc = chan_merge(acc, q);
        }
        double* o = stats_out + tid * 4;
        o[0] = acc.n;
        o[1] = acc.n > 0.0 ? acc.mean : __longlong_as_double(0x7ff8000000000000LL);
        o[2] = acc.m2;
        o[3] = finalize_std(acc);
    }
    if (tid == 0) *ticket = 0u;   // leave the ticket clean for the next launch on this stream
}

__global__ void stats_merge_kernel(const double* __restrict__ parts, int n_parts, int n_feat,
                                   double* __restrict__ out) {
    const int j = threadIdx.x;
    if (j >= n_feat) return;
    Moments acc = {0.0, 0.0, 0.0};
    for (int p = 0; p < n_parts; ++p) {
        const double* q = parts + (size_t(p) * n_feat + j) * 4;
        Moments m = {q[0], (q[0] > 0.0) ? q[1] : 0.0, q[2]};
        acc = chan_merge(acc, m);
    }
    double* o = out + j * 4;
    o[0] = acc.n;
    o[1] = acc.n > 0.0 ? acc.mean : __longlong_as_double(0x7ff8000000000000LL);
    o[2] = acc.m2;
    o[3] = finalize_std(acc);
}

// ------------------------------------------------------------------------------------------
// standardise + pack (preprocess.py:38-41 + train_lightning_ddp.py:45)
//   phase 1: lane = row, loop over columns: coalesced column reads, float64 (x-mean)/std, cast,
//            transposed store into a padded smem tile (odd stride -> conflict-free)
//   phase 2: linear, coalesced row-major write-out of the tile
// ------------------------------------------------------------------------------------------
constexpr int PACK_THREADS = 256;

template <typename T, int F>
__global__ void __launch_bounds__(PACK_THREADS) normalize_pack_kernel(WxCols cols, int64_t n,
                                                                     const double* __restrict__ stats, int impute,
                                                                     float* __restrict__ out,
                                                                     uint32_t* __restrict__ nullmask) {
    constexpr int FP = F | 1;   // odd smem row stride
    __shared__ float tile[PACK_THREADS * FP];
    __shared__ double s_mean[F], s_std[F];
    const int tid = threadIdx.x;
    if (tid < F) {
        s_mean[tid] = stats[tid * 4 + 1];
        s_std[tid] = stats[tid * 4 + 3];
    }
    __syncthreads();
    const int64_t n_tiles = (n + PACK_THREADS - 1) / PACK_THREADS;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const int64_t row0 = t * PACK_THREADS;
        const int64_t row = row0 + tid;
        const bool in = row < n;
        T raw[F];
#pragma unroll
        for (int j = 0; j < F; ++j) raw[j] = in ? static_cast<const T*>(cols.col[j])[row] : T(0);
        uint32_t nm = 0;
#pragma unroll
        for (int j = 0; j < F; ++j) {
            const double x = double(raw[j]);
            bool isnull = isnan(x);
            const uint8_t* vb = cols.valid[j];
            if (vb != nullptr && in) isnull = isnull || !((vb[row >> 3] >> (row & 7)) & 1);
            // exact float64 subtract + divide, then round-to-nearest cast: bit-identical to the oracle
            float v = __double2float_rn((x - s_mean[j]) / s_std[j]);
            if (isnull) {
                nm |= (1u << j);
                v = (impute == WX_IMPUTE_MEAN) ? 0.0f : __int_as_float(0x7fc00000);
            }
            tile[tid * FP + j] = v;
        }
        if (in && nullmask != nullptr) nullmask[row] = nm;
        __syncthreads();
        const int64_t rows_here = (n - row0 < PACK_THREADS) ? (n - row0) : PACK_THREADS;
        const int total = int(rows_here) * F;
        float* dst = out + row0 * F;
#pragma unroll 4
        for (int i = tid; i < total; i += PACK_THREADS) {
            const int r = i / F, c = i - r * F;
            dst[i] = tile[r * FP + c];
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// imputation index list: ordered stream compaction of the set bits of nullmask
// ------------------------------------------------------------------------------------------
constexpr int IDX_THREADS = 256;
constexpr int IDX_RPT = 8;                       // consecutive rows per thread (keeps global order)
constexpr int IDX_TILE = IDX_THREADS * IDX_RPT;

__global__ void __launch_bounds__(IDX_THREADS) idx_count_kernel(const uint32_t* __restrict__ nm, int64_t n,
                                                               unsigned long long* __restrict__ counts) {
    const int64_t base = int64_t(blockIdx.x) * IDX_TILE + int64_t(threadIdx.x) * IDX_RPT;
    int c = 0;
#pragma unroll
    for (int k = 0; k < IDX_RPT; ++k)
        if (base + k < n) c += __popc(nm[base + k]);
    __shared__ int ws[IDX_THREADS / 32];
    int w = c;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = w;
    __syncthreads();
    if (threadIdx.x == 0) {
        int s = 0;
        for (int i = 0; i < IDX_THREADS / 32; ++i) s += ws[i];
        counts[blockIdx.x] = (unsigned long long)s;
    }
}

// single block: exclusive scan of counts[0..nb) in place, total -> *count_out
__global__ void __launch_bounds__(1024) idx_scan_kernel(unsigned long long* __restrict__ counts, int64_t nb,
                                                        int64_t* __restrict__ count_out) {
    __shared__ unsigned long long wsum[32];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0ULL;
    __syncthreads();
    for (int64_t base = 0; base < nb; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const unsigned long long v = (i < nb) ? counts[i] : 0ULL;
        unsigned long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if ((threadIdx.x & 31) >= o) inc += t;
        }
        if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            unsigned long long s = wsum[threadIdx.x];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                unsigned long long t = __shfl_up_sync(0xffffffffu, s, o);
                if (threadIdx.x >= o) s += t;
            }
            wsum[threadIdx.x] = s;   // inclusive over warps
        }
        __syncthreads();
        const unsigned long long warp_off = (threadIdx.x >> 5) ? wsum[(threadIdx.x >> 5) - 1] : 0ULL;
        const unsigned long long c0 = carry;
        if (i < nb) counts[i] = c0 + warp_off + inc - v;   // exclusive
        __syncthreads();
        if (threadIdx.x == 0) carry = c0 + wsum[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) *count_out = (int64_t)carry;
}

__global__ void __launch_bounds__(IDX_THREADS) idx_write_kernel(const uint32_t* __restrict__ nm, int64_t n, int n_feat,
                                                               const unsigned long long* __restrict__ offsets,
                                                               int64_t* __restrict__ idx_out, int64_t capacity) {
    const int64_t base = int64_t(blockIdx.x) * IDX_TILE + int64_t(threadIdx.x) * IDX_RPT;
    uint32_t m[IDX_RPT];
    int c = 0;
#pragma unroll
    for (int k = 0; k < IDX_RPT; ++k) {
        m[k] = (base + k < n) ? nm[base + k] : 0u;
        c += __popc(m[k]);
    }
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if ((threadIdx.x & 31) >= o) inc += t;
    }
    __shared__ int ws[IDX_THREADS / 32];
    if ((threadIdx.x & 31) == 31) ws[threadIdx.x >> 5] = inc;
    __syncthreads();
    int warp_off = 0;
    for (int w = 0; w < int(threadIdx.x >> 5); ++w) warp_off += ws[w];
    int64_t pos = (int64_t)offsets[blockIdx.x] + warp_off + inc - c;
#pragma unroll
    for (int k = 0; k < IDX_RPT; ++k) {
        uint32_t bits = m[k];
        while (bits) {
            const int j = __ffs(bits) - 1;
            bits &= bits - 1;
            if (pos < capacity) idx_out[pos] = (base + k) * n_feat + j;
            ++pos;
        }
    }
}

int fill_cols(const wx_ctx* ctx, const void* const* cols, const uint8_t* const* validity, int dtype, WxCols* wc,
              const char* who) {
    const int f = ctx->desc.n_features;
    WX_CHECK_ARG(cols != nullptr, "%s: cols is NULL", who);
    WX_CHECK_ARG(dtype == WX_F32 || dtype == WX_F64, "%s: dtype=%d is neither WX_F32 nor WX_F64", who, dtype);
    const size_t esz = dtype == WX_F32 ? 4 : 8;
    for (int j = 0; j < WX_MAX_FEATURES; ++j) {
        wc->col[j] = nullptr;
        wc->valid[j] = nullptr;
    }
    for (int j = 0; j < f; ++j) {
        WX_CHECK_ARG(cols[j] != nullptr, "%s: column %d pointer is NULL", who, j);
        WX_CHECK_ARG((reinterpret_cast<uintptr_t>(cols[j]) % esz) == 0, "%s: column %d is not %zu-byte aligned", who, j,
                     esz);
        wc->col[j] = cols[j];
        wc->valid[j] = validity ? validity[j] : nullptr;
    }
    return WX_OK;
}

template <typename T>
void launch_pack(int f, int grid, cudaStream_t st, const WxCols& wc, int64_t n, const double* stats, int impute,
                 float* out, uint32_t* nm) {
#define WX_PACK_CASE(FV) \
    case FV: normalize_pack_kernel<T, FV><<<grid, PACK_THREADS, 0, st>>>(wc, n, stats, impute, out, nm); break;
    switch (f) {
        WX_PACK_CASE(1) WX_PACK_CASE(2) WX_PACK_CASE(3) WX_PACK_CASE(4) WX_PACK_CASE(5) WX_PACK_CASE(6)
        WX_PACK_CASE(7) WX_PACK_CASE(8) WX_PACK_CASE(9) WX_PACK_CASE(10) WX_PACK_CASE(11) WX_PACK_CASE(12)
        WX_PACK_CASE(13) WX_PACK_CASE(14) WX_PACK_CASE(15) WX_PACK_CASE(16) WX_PACK_CASE(17) WX_PACK_CASE(18)
        WX_PACK_CASE(19) WX_PACK_CASE(20) WX_PACK_CASE(21) WX_PACK_CASE(22) WX_PACK_CASE(23) WX_PACK_CASE(24)
        WX_PACK_CASE(25) WX_PACK_CASE(26) WX_PACK_CASE(27) WX_PACK_CASE(28) WX_PACK_CASE(29) WX_PACK_CASE(30)
        WX_PACK_CASE(31) WX_PACK_CASE(32)
        default: break;
    }
#undef WX_PACK_CASE
}

}  // namespace

extern "C" {

int wx_label_encode(wx_ctx* ctx, const int32_t* offsets, const uint8_t* data, const uint8_t* validity, int64_t n,
                    int64_t* labels_out, int32_t* labels32_out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_label_encode: ctx is NULL");
    WX_CHECK_ARG(n >= 0, "wx_label_encode: n=%lld < 0", (long long)n);
    if (n == 0) return WX_OK;
    WX_CHECK_ARG(offsets && labels_out, "wx_label_encode: offsets/labels_out is NULL");
    WX_CHECK_ARG(data != nullptr, "wx_label_encode: data is NULL");
    WX_CUDA(cudaSetDevice(ctx->device));
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = int64_t(ctx->num_sms) * 16;
    if (blocks > cap) blocks = cap;
    label_encode_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(offsets, data, validity, n, labels_out,
                                                                             labels32_out);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

int wx_col_stats(wx_ctx* ctx, const void* const* cols, int dtype, const uint8_t* const* validity, int64_t n,
                 double* stats_out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_col_stats: ctx is NULL");
    WX_CHECK_ARG(n >= 0 && stats_out != nullptr, "wx_col_stats: bad n or NULL stats_out");
    WxCols wc;
    int rc = fill_cols(ctx, cols, validity, dtype, &wc, "wx_col_stats");
    if (rc != WX_OK) return rc;
    WX_CUDA(cudaSetDevice(ctx->device));
    const int f = ctx->desc.n_features;
    const int64_t tile = int64_t(STATS_THREADS) * STATS_G;
    int64_t gx = (n + tile - 1) / tile;
    int64_t cap = (int64_t(ctx->num_sms) * 8 + f - 1) / f;   // ~8 resident CTAs per SM over all columns
    if (cap < 1) cap = 1;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    const size_t need = size_t(f) * gx * 3 * sizeof(double);
    if (WX_SCRATCH_PART_OFF + need > ctx->scratch_bytes) {
        wx_set_error("wx_col_stats: scratch too small");
        return WX_ENOMEM;
    }
    double* parts = reinterpret_cast<double*>(ctx->scratch + WX_SCRATCH_PART_OFF);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(ctx->scratch);
    dim3 grid((unsigned)gx, (unsigned)f);
    if (dtype == WX_F32)
        col_stats_kernel<float><<<grid, STATS_THREADS, 0, (cudaStream_t)stream>>>(wc, f, n, parts, ticket, stats_out);
    else
        col_stats_kernel<double><<<grid, STATS_THREADS, 0, (cudaStream_t)stream>>>(wc, f, n, parts, ticket, stats_out);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

int wx_stats_merge(wx_ctx* ctx, const double* parts, int n_parts, double* out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_stats_merge: ctx is NULL");
    WX_CHECK_ARG(parts && out && n_parts >= 1, "wx_stats_merge: bad arguments");
    WX_CUDA(cudaSetDevice(ctx->device));
    stats_merge_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(parts, n_parts, ctx->desc.n_features, out);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

int wx_normalize_pack(wx_ctx* ctx, const void* const* cols, int dtype, const uint8_t* const* validity, int64_t n,
                      const double* stats, int impute_mode, float* features_out, uint32_t* nullmask_out,
                      void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_normalize_pack: ctx is NULL");
    WX_CHECK_ARG(n >= 0, "wx_normalize_pack: n < 0");
    WX_CHECK_ARG(impute_mode == WX_IMPUTE_NONE || impute_mode == WX_IMPUTE_MEAN, "wx_normalize_pack: impute_mode=%d",
                 impute_mode);
    if (n == 0) return WX_OK;
    WX_CHECK_ARG(stats && features_out, "wx_normalize_pack: stats/features_out is NULL");
    WxCols wc;
    int rc = fill_cols(ctx, cols, validity, dtype, &wc, "wx_normalize_pack");
    if (rc != WX_OK) return rc;
    WX_CUDA(cudaSetDevice(ctx->device));
    const int f = ctx->desc.n_features;
    int64_t tiles = (n + PACK_THREADS - 1) / PACK_THREADS;
    int64_t cap = int64_t(ctx->num_sms) * 8;
    int grid = int(tiles < cap ? tiles : cap);
    if (dtype == WX_F32)
        launch_pack<float>(f, grid, (cudaStream_t)stream, wc, n, stats, impute_mode, features_out, nullmask_out);
    else
        launch_pack<double>(f, grid, (cudaStream_t)stream, wc, n, stats, impute_mode, features_out, nullmask_out);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

int wx_impute_index_list(wx_ctx* ctx, const uint32_t* nullmask, int64_t n, int64_t* idx_out, int64_t capacity,
                         int64_t* count_out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_impute_index_list: ctx is NULL");
    WX_CHECK_ARG(n >= 0 && capacity >= 0 && count_out != nullptr, "wx_impute_index_list: bad arguments");
    WX_CHECK_ARG(capacity == 0 || idx_out != nullptr, "wx_impute_index_list: idx_out is NULL");
    WX_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        WX_CUDA(cudaMemsetAsync(count_out, 0, sizeof(int64_t), st));
        return WX_OK;
    }
    WX_CHECK_ARG(nullmask != nullptr, "wx_impute_index_list: nullmask is NULL");
    const int64_t nb = (n + IDX_TILE - 1) / IDX_TILE;
    if (WX_SCRATCH_PART_OFF + size_t(nb) * 8 > ctx->scratch_bytes) {
        wx_set_error("wx_impute_index_list: n=%lld needs more than the %zu-byte scratch", (long long)n,
                     ctx->scratch_bytes);
        return WX_ENOMEM;
    }
    unsigned long long* counts = reinterpret_cast<unsigned long long*>(ctx->scratch + WX_SCRATCH_PART_OFF);
    idx_count_kernel<<<(unsigned)nb, IDX_THREADS, 0, st>>>(nullmask, n, counts);
    WX_LAUNCH_CHECK(ctx);
    idx_scan_kernel<<<1, 1024, 0, st>>>(counts, nb, count_out);
    WX_LAUNCH_CHECK(ctx);
    idx_write_kernel<<<(unsigned)nb, IDX_THREADS, 0, st>>>(nullmask, n, ctx->desc.n_features, counts, idx_out, capacity);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

}  // extern "C"
