// K1: label encode, column statistics, standardise + pack, imputation index list.
// Replaces the arithmetic of /root/reference/jobs/preprocess.py:23-41 and the casts at
// /root/reference/jobs/train_lightning_ddp.py:45-46.  All HBM-bound streaming kernels.
#include "common.cuh"
#include "tc_common.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// label encode (preprocess.py:23-25)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int32_t label_of(const int32_t a, const int32_t b, const uint8_t* __restrict__ data,
                                            const uint8_t* __restrict__ valid, int64_t i) {
    if (b - a != 4) return 0;
    if (valid != nullptr && !((valid[i >> 3] >> (i & 7)) & 1)) return 0;
    const uint8_t* s = data + a;
    return (s[0] == 'r' && s[1] == 'a' && s[2] == 'i' && s[3] == 'n') ? 1 : 0;
}

// Each thread encodes 4 consecutive rows: one 16-byte load of offsets (+1 scalar), two 16-byte stores of int64 labels.
__global__ void __launch_bounds__(256) label_encode_kernel(const int32_t* __restrict__ off,
                                                          const uint8_t* __restrict__ data,
                                                          const uint8_t* __restrict__ valid, int64_t n,
                                                          int64_t* __restrict__ out64, int32_t* __restrict__ out32,
                                                          int vec_ok) {
    const int64_t stride = int64_t(gridDim.x) * blockDim.x * 4;
    for (int64_t i0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i0 < n; i0 += stride) {
        if (vec_ok && i0 + 4 <= n) {
            const int4 o = *reinterpret_cast<const int4*>(off + i0);
            const int32_t o4 = off[i0 + 4];
            const int32_t l0 = label_of(o.x, o.y, data, valid, i0), l1 = label_of(o.y, o.z, data, valid, i0 + 1);
            const int32_t l2 = label_of(o.z, o.w, data, valid, i0 + 2), l3 = label_of(o.w, o4, data, valid, i0 + 3);
            *reinterpret_cast<longlong2*>(out64 + i0) = make_longlong2(l0, l1);
            *reinterpret_cast<longlong2*>(out64 + i0 + 2) = make_longlong2(l2, l3);
            if (out32) *reinterpret_cast<int4*>(out32 + i0) = make_int4(l0, l1, l2, l3);
        } else {
            for (int64_t i = i0; i < n && i < i0 + 4; ++i) {
                const int32_t lab = label_of(off[i], off[i + 1], data, valid, i);
                out64[i] = lab;
                if (out32) out32[i] = lab;
            }
        }
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
constexpr int STATS_G = 8;   // independent loads per thread per iteration (scalar path)

// Column reference level: the first present, finite value among the column's first 32 rows, else 0.  Every moment inside
// the kernel is carried RELATIVE to it, so a column sitting at 1e9 +- 1 merges group means with the precision of the
// spread, not of the level (a Chan merge is first-order sensitive to errors in the group means).  Warp-collective.
template <typename T>
__device__ __forceinline__ double column_level(const T* __restrict__ col, const uint8_t* __restrict__ vb, int64_t n) {
    const int lane = threadIdx.x & 31;
    double x = 0.0;
    bool ok = false;
    if (lane < n) {
        x = double(col[lane]);
        ok = (fabs(x) <= 1.7976931348623157e308) && (vb == nullptr || ((vb[lane >> 3] >> (lane & 7)) & 1));
    }
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if (m == 0) return 0.0;
    return __shfl_sync(0xffffffffu, x, __ffs(m) - 1);
}

template <typename T>
struct Vec16;
template <>
struct Vec16<float> {
    typedef float4 type;
    static constexpr int N = 4;
    static __device__ __forceinline__ void unpack(const float4& v, float* x) {
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
    }
    // pivots are float-representable, so a NaN cell is replaced by the pivot BEFORE the conversion (one FSEL) and then
    // contributes exactly zero to both sums
    static __device__ __forceinline__ double snap(double k) { return double(__double2float_rn(k)); }
};
template <>
struct Vec16<double> {
    typedef double2 type;
    static constexpr int N = 2;
    static __device__ __forceinline__ void unpack(const double2& v, double* x) {
        x[0] = v.x; x[1] = v.y;
    }
    static __device__ __forceinline__ double snap(double k) { return k; }
};

template <typename T, bool VEC>
__global__ void __launch_bounds__(STATS_THREADS, VEC ? 4 : 3) col_stats_kernel(WxCols cols, int n_feat, int64_t n,
                                                                  double* __restrict__ parts,
                                                                  unsigned int* __restrict__ ticket,
                                                                  double* __restrict__ stats_out) {
    const int j = blockIdx.y;
    const T* __restrict__ col = static_cast<const T*>(cols.col[j]);
    const uint8_t* __restrict__ vb = cols.valid[j];
    const int tid = threadIdx.x;
    const double K0 = column_level<T>(col, vb, n);
    Moments run = {0.0, 0.0, 0.0};   // mean relative to K0
    if (VEC) {
        // 16-byte loads, 4 independent vectors in flight per thread (bytes in flight are what this kernel lives on);
        // the host takes this path when the column is 16-byte aligned and carries no validity bitmap.
        // Shifted sums about a pivot K that is re-centred every FOLD iterations: 3 fp64 ops per cell instead of the
        // divides of a Chan merge per small group.  K is always a value from (or the mean of) this thread's own recent
        // data, so the cancellation in s2 - s1^2/c is bounded by a deviation that itself contributes to M2: the folded
        // group moments are as accurate as two-pass ones to a few ulp, trends and outliers included.
        typedef typename Vec16<T>::type V;
        constexpr int VN = Vec16<T>::N, VG = 4, FOLD = 8;
        const V* __restrict__ colv = reinterpret_cast<const V*>(col);
        const int64_t nv = n / VN;
        const int64_t tilev = int64_t(STATS_THREADS) * VG;
        double K = 0.0, k_rel = 0.0, s1a = 0.0, s1b = 0.0, s2a = 0.0, s2b = 0.0;
        T Kt = T(0);
        int cnt = 0, it = 0;
        bool have_k = false;
        auto fold = [&]() {
            if (cnt > 0) {
                const double s1 = s1a + s1b, s2 = s2a + s2b, c = double(cnt);
                const double dm = s1 / c;
                Moments g;
                g.n = c;
                g.mean = k_rel + dm;
                g.m2 = fmax(s2 - s1 * dm, 0.0);
                run = chan_merge(run, g);
                const double k_new = Vec16<T>::snap(K0 + g.mean);
                if (fabs(k_new) <= 1.7976931348623157e308) {   // keep the old pivot if the data holds infinities
                    K = k_new;
                    Kt = T(K);
                    k_rel = K - K0;
                }
            }
            s1a = s1b = s2a = s2b = 0.0;
            cnt = 0;
            it = 0;
        };
        auto take = [&](const V& v) {
            T e[VN];
            Vec16<T>::unpack(v, e);
            if (!have_k) {
#pragma unroll
                for (int q = 0; q < VN; ++q)
                    if (!have_k && fabs(double(e[q])) <= 1.7976931348623157e308) {
                        K = double(e[q]);
                        Kt = e[q];
                        k_rel = K - K0;
                        have_k = true;
                    }
            }
#pragma unroll
            for (int q = 0; q < VN; ++q) {
                const bool ok = e[q] == e[q];
                cnt += int(ok);
                const T c = ok ? e[q] : Kt;
                const double d = double(c) - K;
                if (q & 1) {
                    s1b += d;
                    s2b = fma(d, d, s2b);
                } else {
                    s1a += d;
                    s2a = fma(d, d, s2a);
                }
            }
        };
        for (int64_t base = int64_t(blockIdx.x) * tilev; base < nv; base += int64_t(gridDim.x) * tilev) {
            if (base + tilev <= nv) {
                V v[VG];
#pragma unroll
                for (int k = 0; k < VG; ++k) v[k] = colv[base + tid + int64_t(k) * STATS_THREADS];
#pragma unroll
                for (int k = 0; k < VG; ++k) take(v[k]);
            } else {
#pragma unroll 1
                for (int k = 0; k < VG; ++k) {
                    const int64_t i = base + tid + int64_t(k) * STATS_THREADS;
                    if (i < nv) take(colv[i]);
                }
            }
            if (++it == FOLD) fold();
        }
        fold();
        // scalar remainder (n not a multiple of the vector width): block 0, thread 0
        if (blockIdx.x == 0 && tid == 0) {
            for (int64_t i = nv * VN; i < n; ++i) {
                const double xv = double(col[i]);
                if (!isnan(xv)) {
                    Moments g = {1.0, xv - K0, 0.0};
                    run = chan_merge(run, g);
                }
            }
        }
    } else {
        const int64_t tile = int64_t(STATS_THREADS) * STATS_G;
        for (int64_t base = int64_t(blockIdx.x) * tile; base < n; base += int64_t(gridDim.x) * tile) {
            double x[STATS_G];
            bool ok[STATS_G];
#pragma unroll
            for (int k = 0; k < STATS_G; ++k) {
                const int64_t i = base + tid + int64_t(k) * STATS_THREADS;
                ok[k] = i < n;
                x[k] = ok[k] ? double(col[i]) - K0 : 0.0;
                if (ok[k] && vb != nullptr) ok[k] = (vb[i >> 3] >> (i & 7)) & 1;
                ok[k] = ok[k] && !isnan(x[k]);
            }
            double c = 0.0, sum = 0.0;
#pragma unroll
            for (int k = 0; k < STATS_G; ++k)
                if (ok[k]) {
                    c += 1.0;
                    sum += x[k];
                }
            if (c > 0.0) {
                Moments g;
                g.n = c;
                g.mean = sum / c;
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
        double* p = parts + (size_t(j) * gridDim.x + blockIdx.x) * 4;
        p[0] = b.n;
        p[1] = b.mean;
        p[2] = b.m2;
        p[3] = K0;
        __threadfence();
        const unsigned int total = gridDim.x * gridDim.y;
        is_last = (atomicAdd(ticket, 1u) == total - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    if (tid < n_feat) {
        const volatile double* p = parts + size_t(tid) * gridDim.x * 4;
        Moments acc = {0.0, 0.0, 0.0};
        for (unsigned int b = 0; b < gridDim.x; ++b) {
            Moments q = {p[b * 4 + 0], p[b * 4 + 1], p[b * 4 + 2]};
            acc = chan_merge(acc, q);
        }
        double* o = stats_out + tid * 4;
        o[0] = acc.n;
        o[1] = acc.n > 0.0 ? p[3] + acc.mean : __longlong_as_double(0x7ff8000000000000LL);
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
// One cell of preprocess.py:41 + train_lightning_ddp.py:45: float32((x - mean) / std) with the subtract and divide in
// float64, bit-identical to the oracle -- but the divide is replaced by a multiply with 1/std whenever that provably
// rounds to the same fp32 value.  q = fl(d * fl(1/std)) is within 2 fp64 ulps of the true quotient; the cast to fp32
// drops the low 29 mantissa bits, so unless those sit within 16 ulps of the rounding midpoint (about one cell in 2^24)
// the cast of q IS the cast of the quotient.  Non-zero results below the fp32 normal range (a wider discarded field)
// and columns whose 1/std is not a full-precision double (`force`) take the real divide.  NaN in gives NaN out either
// way, and an exact zero is exact.  This keeps the kernel on the HBM roofline instead of the fp64 divide sequence.
__device__ __forceinline__ float zscore_cell(double x, double mean, double sd, double rstd, bool force) {
    const double d = x - mean;
    const double q = d * rstd;
    float v = __double2float_rn(q);
    const uint32_t lo = uint32_t(__double2loint(q)), hq = uint32_t(__double2hiint(q)) & 0x7FFFFFFFu;
    const bool near_mid = ((lo & 0x1FFFFFFFu) - 0x0FFFFFF0u) <= 0x20u;
    const bool tiny = (hq - 1u) < 0x380FFFFFu;   // 0 < |q| < 2^-126
    if (near_mid || tiny || force) v = __double2float_rn(d / sd);
    return v;
}

constexpr int PACK_THREADS = 256;

// lane = row: coalesced column reads into registers, the cell math, then the row goes to a shared-memory tile.
//   F % 4 == 0: the tile is the dense row-major image of the output (16-byte STS, conflict-free when F/4 is odd) and
//               leaves through ONE TMA bulk store issued by thread 0; two tile buffers, so the next tile's loads and
//               math overlap the store -- one __syncthreads per tile and no per-cell store instructions at all.
//   otherwise : odd row stride (conflict-free scalar STS) and a cooperative coalesced copy-out.
template <typename T, int F>
__global__ void __launch_bounds__(PACK_THREADS, (sizeof(T) == 8 ? 3 : 4)) normalize_pack_kernel(WxCols cols, int64_t n,
                                                                        const double* __restrict__ stats, int impute,
                                                                        float* __restrict__ out,
                                                                        uint32_t* __restrict__ nullmask,
                                                                        int out_aligned16) {
    constexpr bool DENSE = (F % 4 == 0);
    constexpr int TS = DENSE ? F : (F | 1);
    constexpr int NBUF = DENSE ? 2 : 1;
    extern __shared__ __align__(128) float pack_tile[];   // [NBUF][PACK_THREADS * TS]
    __shared__ double s_mean[F], s_std[F], s_rstd[F];
    __shared__ int s_force;
    const int tid = threadIdx.x;
    if (tid == 0) s_force = 0;
    __syncthreads();
    if (tid < F) {
        const double sd = stats[tid * 4 + 3];
        s_mean[tid] = stats[tid * 4 + 1];
        s_std[tid] = sd;
        const bool full = fabs(sd) > 1e-290 && fabs(sd) < 1e290;   // 1/std normal and correctly rounded
        s_rstd[tid] = full ? 1.0 / sd : 0.0;
        if (!full) s_force = 1;   // benign race: every writer stores 1
    }
    __syncthreads();
    const bool force = s_force != 0;
    const bool bulk = DENSE && out_aligned16;
    const float null_val = (impute == WX_IMPUTE_MEAN) ? 0.0f : __int_as_float(0x7fc00000);
    const int64_t n_tiles = (n + PACK_THREADS - 1) / PACK_THREADS;
    int k = 0;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
        const int64_t row0 = t * PACK_THREADS;
        const int64_t row = row0 + tid;
        const bool in = row < n;
        float* tb = pack_tile + ((bulk && NBUF == 2) ? (k & 1) : 0) * (PACK_THREADS * TS);
        T raw[F];
#pragma unroll
        for (int j = 0; j < F; ++j) raw[j] = in ? static_cast<const T*>(cols.col[j])[row] : T(0);
        uint32_t nm = 0;
#pragma unroll
        for (int j0 = 0; j0 < F; j0 += 4) {
            float vq[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int j = j0 + e;
                if (j < F) {
                    bool isnull = raw[j] != raw[j];
                    const uint8_t* vb = cols.valid[j];
                    if (vb != nullptr && in) isnull = isnull || !((vb[row >> 3] >> (row & 7)) & 1);
                    const float v = zscore_cell(double(raw[j]), s_mean[j], s_std[j], s_rstd[j], force);
                    if (isnull) nm |= (1u << j);
                    vq[e] = isnull ? null_val : v;
                    if (!DENSE) tb[tid * TS + j] = vq[e];
                }
            }
            if (DENSE) *reinterpret_cast<float4*>(tb + tid * TS + j0) = make_float4(vq[0], vq[1], vq[2], vq[3]);
        }
        if (in && nullmask != nullptr) nullmask[row] = nm;
        const int64_t rows_here = (n - row0 < PACK_THREADS) ? (n - row0) : PACK_THREADS;
        float* dst = out + row0 * F;
        if (bulk) {
            // make this thread's tile writes visible to the async proxy; thread 0 additionally makes sure the store of
            // the previous tile has finished READING the other buffer before anyone (after the barrier) refills it
            wxtc::fence_proxy_async();
            if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncthreads();
            if (tid == 0) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst),
                             "r"(wxtc::smem_u32(tb)), "r"(uint32_t(rows_here) * F * 4u)
                             : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        } else {
            __syncthreads();
            const int total = int(rows_here) * F;
#pragma unroll 4
            for (int i = tid; i < total; i += PACK_THREADS) {
                const int r = i / F, c = i - r * F;
                dst[i] = tb[r * TS + c];
            }
            __syncthreads();
        }
    }
    if (bulk && tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
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

// ---- columnar fp64 z-score: the `*_norm` columns exactly as preprocess.py:38-41 writes them ------------------------
// out_j[i] = ((double)col_j[i] - mean_j) / std_val_j with IEEE float64 subtraction and DIVISION (not a reciprocal
// multiply): bit-identical to the float64 expression Spark evaluates.  A clear validity bit gives NaN (SQL null).
// HBM-bound: 8 (4) bytes in, 8 bytes out per cell; grid.y = column, each thread owns pairs of rows (16-byte stores).
struct WxOutCols {
    double* col[WX_MAX_FEATURES];
};

template <typename T>
__global__ void __launch_bounds__(256) zscore_f64_kernel(const WxCols in, const WxOutCols out, int64_t n,
                                                          const double* __restrict__ stats) {
    const int j = blockIdx.y;
    const T* __restrict__ x = static_cast<const T*>(in.col[j]);
    const uint8_t* __restrict__ valid = in.valid[j];
    double* __restrict__ o = out.col[j];
    const double mean = stats[j * 4 + 1], sd = stats[j * 4 + 3];
    const bool vec = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(o)) & 15) == 0;
    const int64_t npairs = n >> 1;
    const int64_t stride = int64_t(gridDim.x) * blockDim.x;
    if (vec) {
        for (int64_t pidx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; pidx < npairs; pidx += stride) {
            const int64_t i = pidx * 2;
            double a, b;
            if (sizeof(T) == 8) {
                const double2 v = *reinterpret_cast<const double2*>(x + i);
                a = v.x; b = v.y;
            } else {
                const float2 v = *reinterpret_cast<const float2*>(x + i);
                a = double(v.x); b = double(v.y);
            }
            a = (a - mean) / sd;
            b = (b - mean) / sd;
            if (valid != nullptr) {
                const uint8_t bits = valid[i >> 3] >> (i & 7);         // i is even: both bits sit in the same byte
                if (!(bits & 1)) a = __longlong_as_double(0x7ff8000000000000ll);
                if (!(bits & 2)) b = __longlong_as_double(0x7ff8000000000000ll);
            }
            *reinterpret_cast<double2*>(o + i) = make_double2(a, b);
        }
    }
    // unaligned columns: scalar over everything; aligned: only the odd last row
    const int64_t first = vec ? (npairs << 1) : 0;
    for (int64_t i = first + int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        double a = (double(x[i]) - mean) / sd;
        if (valid != nullptr && !((valid[i >> 3] >> (i & 7)) & 1)) a = __longlong_as_double(0x7ff8000000000000ll);
        o[i] = a;
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

template <typename T, int F>
int launch_pack_f(const wx_ctx* ctx, cudaStream_t st, const WxCols& wc, int64_t n, const double* stats,
                  int impute, float* out, uint32_t* nm) {
    const int64_t tiles = (n + PACK_THREADS - 1) / PACK_THREADS;
    const int64_t cap = int64_t(ctx->num_sms) * (sizeof(T) == 8 ? 3 : 4);   // one resident wave, persistent tiles
    const int grid = int(tiles < cap ? tiles : cap);
    const int aligned = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    constexpr bool dense = (F % 4 == 0);
    constexpr size_t smem = size_t(dense ? 2 : 1) * PACK_THREADS * (dense ? F : (F | 1)) * sizeof(float);
    if (smem > 48 * 1024) {
        static bool attr_set[64] = {};
        if (!attr_set[ctx->device & 63]) {
            WX_CUDA(cudaFuncSetAttribute(normalize_pack_kernel<T, F>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         int(smem)));
            attr_set[ctx->device & 63] = true;
        }
    }
    normalize_pack_kernel<T, F><<<grid, PACK_THREADS, smem, st>>>(wc, n, stats, impute, out, nm, aligned);
    return WX_OK;
}

template <typename T>
int launch_pack(const wx_ctx* ctx, int f, cudaStream_t st, const WxCols& wc, int64_t n,
                const double* stats, int impute, float* out, uint32_t* nm) {
#define WX_PACK_CASE(FV) \
    case FV: return launch_pack_f<T, FV>(ctx, st, wc, n, stats, impute, out, nm);
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
    return WX_OK;
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
    int64_t blocks = (n + 1023) / 1024;
    const int64_t cap = int64_t(ctx->num_sms) * 16;
    if (blocks > cap) blocks = cap;
    const uintptr_t al = reinterpret_cast<uintptr_t>(offsets) | reinterpret_cast<uintptr_t>(labels_out) |
                         (labels32_out ? reinterpret_cast<uintptr_t>(labels32_out) : 0);
    label_encode_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(offsets, data, validity, n, labels_out,
                                                                             labels32_out, (al & 15) == 0);
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
    int64_t cap = (int64_t(ctx->num_sms) * 8 + f - 1) / f;   // ~8 resident CTAs per SM over all columns (256 B in flight per thread)
    if (cap < 1) cap = 1;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    const size_t need = size_t(f) * gx * 4 * sizeof(double);
    if (WX_SCRATCH_PART_OFF + need > ctx->scratch_bytes) {
        wx_set_error("wx_col_stats: scratch too small");
        return WX_ENOMEM;
    }
    double* parts = reinterpret_cast<double*>(ctx->scratch + WX_SCRATCH_PART_OFF);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(ctx->scratch);
    dim3 grid((unsigned)gx, (unsigned)f);
    bool vec = validity == nullptr;
    for (int j = 0; j < f && vec; ++j) vec = (reinterpret_cast<uintptr_t>(cols[j]) & 15) == 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == WX_F32) {
        if (vec) col_stats_kernel<float, true><<<grid, STATS_THREADS, 0, st>>>(wc, f, n, parts, ticket, stats_out);
        else col_stats_kernel<float, false><<<grid, STATS_THREADS, 0, st>>>(wc, f, n, parts, ticket, stats_out);
    } else {
        if (vec) col_stats_kernel<double, true><<<grid, STATS_THREADS, 0, st>>>(wc, f, n, parts, ticket, stats_out);
        else col_stats_kernel<double, false><<<grid, STATS_THREADS, 0, st>>>(wc, f, n, parts, ticket, stats_out);
    }
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
    if (dtype == WX_F32)
        rc = launch_pack<float>(ctx, f, (cudaStream_t)stream, wc, n, stats, impute_mode, features_out,
                                nullmask_out);
    else
        rc = launch_pack<double>(ctx, f, (cudaStream_t)stream, wc, n, stats, impute_mode, features_out,
                                 nullmask_out);
    if (rc != WX_OK) return rc;
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

int wx_zscore_f64(wx_ctx* ctx, const void* const* cols, int dtype, const uint8_t* const* validity, int64_t n,
                  const double* stats, double* const* out_cols, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_zscore_f64: ctx is NULL");
    WX_CHECK_ARG(n >= 0, "wx_zscore_f64: n < 0");
    if (n == 0) return WX_OK;
    WX_CHECK_ARG(stats && out_cols, "wx_zscore_f64: stats/out_cols is NULL");
    WxCols wc;
    int rc = fill_cols(ctx, cols, validity, dtype, &wc, "wx_zscore_f64");
    if (rc != WX_OK) return rc;
    const int f = ctx->desc.n_features;
    WxOutCols oc;
    for (int j = 0; j < WX_MAX_FEATURES; ++j) oc.col[j] = nullptr;
    for (int j = 0; j < f; ++j) {
        WX_CHECK_ARG(out_cols[j] != nullptr && (reinterpret_cast<uintptr_t>(out_cols[j]) & 7) == 0,
                     "wx_zscore_f64: output column %d is NULL or not 8-byte aligned", j);
        oc.col[j] = out_cols[j];
    }
    WX_CUDA(cudaSetDevice(ctx->device));
    int64_t gx = ((n >> 1) + 255) / 256;
    const int64_t cap = (int64_t(ctx->num_sms) * 8 + f - 1) / f;      // ~8 resident CTAs per SM over all columns
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    dim3 grid((unsigned)gx, (unsigned)f);
    if (dtype == WX_F32) zscore_f64_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(wc, oc, n, stats);
    else zscore_f64_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>(wc, oc, n, stats);
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
