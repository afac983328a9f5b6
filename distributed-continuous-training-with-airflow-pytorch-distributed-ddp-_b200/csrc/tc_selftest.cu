// Bring-up kernels for the tcgen05 path: the two GEMM shapes of the tensor-core step kernel in isolation, so the
// descriptor encodings, swizzled layouts and TMEM read-out can be checked against a host matmul.
//   selftest 1:  Z[128][64]  = Xaug[128][32] * Waug[64][32]^T      (A, B K-major; M=128, N=64, K=32; 3-term tf32 split)
//   selftest 2:  G[64][32]   = DH[128][64]^T * Xaug[128][32]       (A, B MN-major; M=64, N=32, K=128; 3-term split)
#include "common.cuh"
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
        store_sw(s.xhi, t, c, hi);
        store_sw(s.xlo, t, c, lo);
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
                const uint64_t ad = make_smem_desc(smem_u32(at[term]) + k * 32, 16, 1024);
                const uint64_t bd = make_smem_desc(smem_u32(bt[term]) + k * 32, 16, 1024);
                umma_tf32(d1, ad, bd, id1, acc);
                acc = true;
            }
        }
        // ---- GEMM2: M=64, N=32, K=128 (16 k-steps of 8), operands MN-major ----
        const uint32_t id2 = make_idesc_tf32(64, 32, true, true);
        const float* a2[3] = {s.ahi, s.alo, s.ahi};
        const float* b2[3] = {s.xmlo, s.xmhi, s.xmhi};
        acc = false;
        for (int term = 3 - terms; term < 3; ++term) {
            for (int k = 0; k < 16; ++k) {
                const uint64_t ad = make_smem_desc(smem_u32(a2[term]) + k * 2048, 512, 1024, LAYOUT_SW128_B32);
                const uint64_t bd = make_smem_desc(smem_u32(b2[term]) + k * 1024, 512, 512, LAYOUT_SW128_B32);
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

extern "C" int wx_tc_selftest(wx_ctx* ctx, const float* x, const float* w, const float* dh, int terms, float* z, float* g,
                              int* status, void* stream) {
    WX_CHECK_ARG(ctx != nullptr, "wx_tc_selftest: ctx is NULL");
    WX_CHECK_ARG(x && w && dh && z && g && status, "wx_tc_selftest: NULL pointer");
    WX_CHECK_ARG(terms >= 1 && terms <= 3, "wx_tc_selftest: terms=%d", terms);
    WX_CUDA(cudaSetDevice(ctx->device));
    const int smem = int(sizeof(SelfSmem)) + 1024;
    WX_CUDA(cudaFuncSetAttribute(tc_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tc_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(x, w, dh, terms, z, g, status);
    WX_LAUNCH_CHECK(ctx);
    return WX_OK;
}
