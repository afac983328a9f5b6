// Bring-up kernels for the tcgen05 path: the two GEMM shapes of the tensor-core step kernel in isolation, so the
// descriptor encodings, swizzled layouts and TMEM read-out can be checked against a host matmul.
//   selftest 1:  Z[128][64]  = Xaug[128][32] * Waug[64][32]^T      (A, B K-major; M=128, N=64, K=32; 3-term tf32 split)
//   selftest 2:  G[64][32]   = DH[128][64]^T * Xaug[128][32]       (A, B MN-major; M=64, N=32, K=128; 3-term split)
#include "common.cuh"
#include "wxddp_dev.h"
#include "tc_common.cuh"

namespace {
using namespace wxtc;

struct __align__(1024) SelfSmem {
    float xhi[128 * 32];     // [row][32] SW128 (K-major A of GEMM1)
    float xlo[128 * 32];
    float xmhi[128 * 32];    // [row][32] SW128_BASE32B (MN-major B of GEMM2)
    float xmlo[128 * 32];
    float whi[64 * 32];      // [n][32] swizzled
    float wlo[64 * 32];
    float ahi[16 * 2 * 256]; // GEMM2 A: 16 k-atoms x 2 m-atoms x 1024 B
    float alo[16 * 2 * 256];
    uint64_t bar;
    uint32_t tmem_base;
    int err;
};

__device__ __forceinline__ void store_sw(float* tile, uint32_t row, uint32_t col, float v) {
    *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(tile) + sw128_offset(row, col)) = v;
}

__global__ void __launch_bounds__(128) tc_selftest_kernel(const float* __restrict__ x /*[128][32]*/,
                                                          const float* __restrict__ w /*[64][32]*/,
                                                          const float* __restrict__ dh /*[128][64]*/, int terms,
                                                          float* __restrict__ z /*[128][64]*/,
                                                          float* __restrict__ g /*[64][32]*/, int* __restrict__ status) {
    extern __shared__ __align__(1024) unsigned char raw[];
    SelfSmem& s = *reinterpret_cast<SelfSmem*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    // terms 11..13: variant with ONE unswizzled X buffer (8 x 16-byte core matrices, column groups 128 B apart, 8-row
    // groups 1024 B apart) read K-major by GEMM1 and MN-major by GEMM2
    const bool shared_x = terms > 10;
    if (shared_x) terms -= 10;
    if (t == 0) {
        mbar_init(&s.bar, 1);
        fence_mbar_init();
        s.err = 0;
    }
    if (warp == 0) tmem_alloc(&s.tmem_base, 128);
    // operands: thread t = row t of Xaug (and row t of W for t < 64)
    for (int c = 0; c < 32; ++c) {
        float hi, lo;
        split_tf32(x[t * 32 + c], hi, lo);
        if (shared_x) {
            const uint32_t off = (t >> 3) * 1024 + (c >> 2) * 128 + (t & 7) * 16 + (c & 3) * 4;
            *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(s.xhi) + off) = hi;
            *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(s.xlo) + off) = lo;
        } else {
            store_sw(s.xhi, t, c, hi);
            store_sw(s.xlo, t, c, lo);
        }
        *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(s.xmhi) + mn32_offset(c, t, 512, 512)) = hi;
        *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(s.xmlo) + mn32_offset(c, t, 512, 512)) = lo;
        if (t < 64) {
            split_tf32(w[t * 32 + c], hi, lo);
            store_sw(s.whi, t, c, hi);
            store_sw(s.wlo, t, c, lo);
        }
    }
    // GEMM2 A operand, MN-major (BASE32B): element (m = hidden j, k = row t); atoms (ki, mi) at (ki*2 + mi)*512 B
    for (int j = 0; j < 64; ++j) {
        float hi, lo;
        split_tf32(dh[t * 64 + j], hi, lo);
        const uint32_t off = mn32_offset(j, t, 512, 1024);
        *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(s.ahi) + off) = hi;
        *reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(s.alo) + off) = lo;
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = s.tmem_base;
    const uint32_t d1 = tm;          // columns [0,64)
    const uint32_t d2 = tm + 64;     // columns [64,96)
    if (t == 0) {
        // ---- GEMM1: M=128, N=64, K=32 (4 k-steps of 8), operands K-major ----
        const uint32_t id1 = make_idesc_tf32(128, 64, false, false);
        const float* at[3] = {s.xhi, s.xlo, s.xhi};
        const float* bt[3] = {s.wlo, s.whi, s.whi};
        bool acc = false;
        for (int term = 3 - terms; term < 3; ++term) {
            for (int k = 0; k < 4; ++k) {
                const uint64_t ad = shared_x ? make_smem_desc(smem_u32(at[term]) + k * 256, 128, 1024, 0)
                                             : make_smem_desc(smem_u32(at[term]) + k * 32, 16, 1024);
                const uint64_t bd = make_smem_desc(smem_u32(bt[term]) + k * 32, 16, 1024);
                umma_tf32(d1, ad, bd, id1, acc);
                acc = true;
            }
        }
        // ---- GEMM2: M=64, N=32, K=128 (16 k-steps of 8), operands MN-major ----
        const uint32_t id2 = make_idesc_tf32(64, 32, true, true);
        const float* a2[3] = {s.ahi, s.alo, s.ahi};
        const float* b2[3] = {shared_x ? s.xlo : s.xmlo, shared_x ? s.xhi : s.xmhi, shared_x ? s.xhi : s.xmhi};
        acc = false;
        for (int term = 3 - terms; term < 3; ++term) {
            for (int k = 0; k < 16; ++k) {
                const uint64_t ad = make_smem_desc(smem_u32(a2[term]) + k * 2048, 512, 1024, LAYOUT_SW128_B32);
                const uint64_t bd = shared_x ? make_smem_desc(smem_u32(b2[term]) + k * 1024, 1024, 128, 0)
                                             : make_smem_desc(smem_u32(b2[term]) + k * 1024, 512, 512, LAYOUT_SW128_B32);
                umma_tf32(d2, ad, bd, id2, acc);
                acc = true;
            }
        }
        umma_commit(&s.bar);
    }
    __syncwarp();
    if (!mbar_wait(&s.bar, 0)) s.err = 1;
    tc_fence_after();
    // read-out: warp w owns TMEM lanes 32w .. 32w+31
    const uint32_t lane_base = uint32_t(warp * 32) << 16;
    for (int c = 0; c < 64; c += 16) {
        float v[16];
        tmem_ld16(d1 + lane_base + c, v);
        tmem_ld_wait();
        for (int i = 0; i < 16; ++i) z[t * 64 + c + i] = v[i];
    }
    for (int c = 0; c < 32; c += 16) {
        float v[16];
        tmem_ld16(d2 + lane_base + c, v);
        tmem_ld_wait();
        if (lane < 16)       // M=64 accumulators live in lanes 0-15 of each 32-lane sub-partition: m = 16*warp + lane
            for (int i = 0; i < 16; ++i) g[(warp * 16 + lane) * 32 + c + i] = v[i];
    }
    tc_fence_before();
    __syncthreads();
    if (t == 0) *status = s.err;
    if (warp == 0) tmem_dealloc(tm, 128);
}

}  // namespace

extern "C" int wxdev_tc_selftest(wx_ctx* ctx, const float* x, const float* w, const float* dh, int terms, float* z, float* g,
                              int* status, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wxdev_tc_selftest: ctx is NULL");
    WX_CHECK_ARG(x && w && dh && z && g && status, "wxdev_tc_selftest: NULL pointer");
    WX_CHECK_ARG((terms >= 1 && terms <= 3) || (terms >= 11 && terms <= 13), "wxdev_tc_selftest: terms=%d", terms);
    WX_CUDA(cudaSetDevice(ctx->device));
    const int smem = int(sizeof(SelfSmem)) + 1024;
    WX_CUDA(cudaFuncSetAttribute(tc_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tc_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(x, w, dh, terms, z, g, status);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}

// ---- MMA issue/throughput micro-benchmark (developer aid): cycles per tcgen05.mma for the operand layouts in play ----
namespace {
using namespace wxtc;

__global__ void __launch_bounds__(128) tc_mma_bench_kernel(int variant, int n_dim, int reps, long long* __restrict__ out) {
    extern __shared__ __align__(1024) unsigned char raw[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int t = threadIdx.x, warp = t >> 5;
    for (int i = t; i < 96 * 1024 / 4; i += 128) reinterpret_cast<float*>(raw)[i] = 0.0f;
    if (t == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc(&tmem_base, 128);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (variant >= 8 && warp == 0) {
        // warp-collective issue: whole warp runs the loop, one elected lane issues each MMA
        const uint32_t base = smem_u32(raw);
        const long long t0 = clock64();
        if (variant == 8) {
            const uint32_t id = make_idesc_tf32(128, n_dim, false, false);
            const uint64_t ad = make_smem_desc(base, 16, 1024), bd = make_smem_desc(base + 32768, 16, 1024);
            for (int r = 0; r < reps; ++r) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_tf32_warp(tmem_base, ad + uint64_t(2 * k), bd + uint64_t(2 * k), id, true);
            }
        } else {
            const uint32_t id = make_idesc_tf32(64, n_dim, true, true);
            const uint64_t ad = make_smem_desc(base, 512, 1024, LAYOUT_SW128_B32), bd = make_smem_desc(base + 65536, 512, 512, LAYOUT_SW128_B32);
            for (int r = 0; r < reps; ++r) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_tf32_warp(tmem_base, ad + uint64_t(128 * k), bd + uint64_t(64 * k), id, true);
            }
        }
        const long long t1 = clock64();
        umma_commit_warp(&bar);
        mbar_wait_hot(&bar, 0);
        const long long t2 = clock64();
        if (blockIdx.x == 0 && t == 0) {
            out[0] = t1 - t0;
            out[1] = t2 - t0;
        }
    } else if (variant < 8 && t == 0) {
        const uint32_t base = smem_u32(raw);
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            if (variant == 0) {          // GEMM1 as built: K-major SW128, M=128, N=n_dim, 4 k-steps inside one 128-B row
                const uint32_t id = make_idesc_tf32(128, n_dim, false, false);
                for (int k = 0; k < 4; ++k)
                    umma_tf32(tmem_base, make_smem_desc(base + k * 32, 16, 1024), make_smem_desc(base + 32768 + k * 32, 16, 1024), id, true);
            } else if (variant == 1) {   // GEMM1 candidate: K-major, no swizzle, one contiguous [rows][8] block per k-step
                const uint32_t id = make_idesc_tf32(128, n_dim, false, false);
                for (int k = 0; k < 4; ++k)
                    umma_tf32(tmem_base, make_smem_desc(base + k * 4096, 128, 256, 0), make_smem_desc(base + 32768 + k * 2048, 128, 256, 0), id, true);
            } else if (variant == 2) {   // GEMM2 as built: MN-major SW128_B32, M=64, N=n_dim
                const uint32_t id = make_idesc_tf32(64, n_dim, true, true);
                for (int k = 0; k < 4; ++k)
                    umma_tf32(tmem_base, make_smem_desc(base + k * 2048, 512, 1024, LAYOUT_SW128_B32),
                              make_smem_desc(base + 65536 + k * 1024, 512, 512, LAYOUT_SW128_B32), id, true);
            } else if (variant == 4) {   // GEMM1 layout, four independent accumulators (columns 0,32,64,96), N <= 32
                const uint32_t id = make_idesc_tf32(128, n_dim, false, false);
                const uint64_t ad = make_smem_desc(base, 16, 1024), bd = make_smem_desc(base + 32768, 16, 1024);
                for (int k = 0; k < 4; ++k) umma_tf32(tmem_base + 32 * k, ad + uint64_t(2 * k), bd + uint64_t(2 * k), id, true);
            } else if (variant == 5) {   // GEMM1 layout, one accumulator, descriptors hoisted (pure issue cost of a dependent chain)
                const uint32_t id = make_idesc_tf32(128, n_dim, false, false);
                const uint64_t ad = make_smem_desc(base, 16, 1024), bd = make_smem_desc(base + 32768, 16, 1024);
                for (int k = 0; k < 4; ++k) umma_tf32(tmem_base, ad + uint64_t(2 * k), bd + uint64_t(2 * k), id, true);
            } else if (variant == 6) {   // GEMM2 layout, two independent accumulators
                const uint32_t id = make_idesc_tf32(64, n_dim, true, true);
                const uint64_t ad = make_smem_desc(base, 512, 1024, LAYOUT_SW128_B32), bd = make_smem_desc(base + 65536, 512, 512, LAYOUT_SW128_B32);
                for (int k = 0; k < 4; ++k) umma_tf32(tmem_base + 32 * (k & 1), ad + uint64_t(128 * k), bd + uint64_t(64 * k), id, true);
            } else {                     // GEMM2 candidate: K-major no swizzle, M=64, N=n_dim
                const uint32_t id = make_idesc_tf32(64, n_dim, false, false);
                for (int k = 0; k < 4; ++k)
                    umma_tf32(tmem_base, make_smem_desc(base + k * 2048, 128, 256, 0), make_smem_desc(base + 65536 + k * 1024, 128, 256, 0), id, true);
            }
        }
        const long long t1 = clock64();
        umma_commit(&bar);
        mbar_wait_hot(&bar, 0);
        const long long t2 = clock64();
        if (blockIdx.x == 0) {
            out[0] = t1 - t0;
            out[1] = t2 - t0;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 128);
}
}  // namespace

extern "C" int wxdev_tc_mma_bench(wx_ctx* ctx, int variant, int n_dim, int reps, long long* out, void* stream) {
    WX_CHECK_ARG(ctx != nullptr && out != nullptr, "wxdev_tc_mma_bench: NULL argument");
    WX_CUDA(cudaSetDevice(ctx->device));
    const int smem = 97 * 1024;
    WX_CUDA(cudaFuncSetAttribute(tc_mma_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tc_mma_bench_kernel<<<ctx->num_sms, 128, smem, (cudaStream_t)stream>>>(variant, n_dim, reps, out);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}
