// Blackwell (sm_100a) primitives used by the tensor-core step kernel: mbarrier, TMA bulk copy, TMEM
// allocation / loads, tcgen05.mma (kind::tf32) and the shared-memory / instruction descriptors.
// Inline PTX only -- no CUTLASS dependency.  Bit layouts follow the tcgen05 descriptor format
// (start address, LBO, SBO in 16-byte units; version 1; layout type in bits 61..63).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace wxtc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// non-blocking test (try_wait may park the thread for a hardware time slice; a scheduler that polls two barriers must not)
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// try_wait with a suspend-time hint: the hardware parks the thread for up to `ns` instead of returning after its default
// time slice, so a waiting role issues one instruction per `ns` rather than a polling loop's worth
__device__ __forceinline__ bool mbar_try_wait_hint(uint64_t* bar, uint32_t parity, uint32_t ns) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(ns)
        : "memory");
    return ok != 0;
}

// Bounded wait: returns false if the barrier did not flip within ~2 s (a protocol bug must not hang the GPU).
// A failed try_wait comes back after a few dozen cycles whatever the hint says; an idle role that kept polling at that
// rate took a quarter of the SM's issue slots from the roles doing the work, hence the sleep between polls.
template <unsigned SLEEP_NS = 128>
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return true;
#pragma unroll 1
    for (uint32_t spin = 0; spin < (1u << 23); ++spin) {
        __nanosleep(SLEEP_NS);
        if (mbar_try_wait(bar, parity)) return true;
    }
    return false;
}

// Tight polling (try_wait itself blocks for a hardware time slice): for roles that own their scheduler slot, e.g. the
// producer and MMA warps of the wide GEMM.
__device__ __forceinline__ bool mbar_wait_spin(uint64_t* bar, uint32_t parity) {
#pragma unroll 1
    for (uint32_t spin = 0; spin < (1u << 28); ++spin)
        if (mbar_try_wait(bar, parity)) return true;
    return false;
}

// Latency-critical variant for the MMA issuers of the step kernel: short sleeps.
__device__ __forceinline__ bool mbar_wait_hot(uint64_t* bar, uint32_t parity) { return mbar_wait<32>(bar, parity); }

// generic-proxy writes (st.shared) -> visible to the async proxy (tcgen05.mma / TMA reads of smem)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMA 1-D bulk copy global -> shared, completion on an mbarrier ---------------------------------------
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ---- 16-byte cp.async (LDGSTS, L2 only) with completion on an mbarrier ------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
// the executing thread's arrive on `bar` happens when all of ITS earlier cp.async operations have landed (.noinc: the
// arrive is one of the barrier's expected arrivals)
__device__ __forceinline__ void cp_async_mbar_arrive(uint64_t* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- TMEM ------------------------------------------------------------------------------------------------
// One full warp calls alloc; the base address (lane 0, first column) lands in *smem_out.
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_out, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_out)), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 16 consecutive 32-bit columns: thread i of the warp gets lane (base_lane + i), columns c..c+15
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Variants without a "memory" clobber, so that ordinary shared-memory loads can be scheduled across them: tmem_ld16_nc
// issues the load, tmem_ld_wait_dep waits for ALL outstanding loads and names the 16 result registers as read-write
// operands -- every use of them is data-dependent on the wait and cannot be scheduled before it.  (volatile asm
// statements keep their order among themselves.)
__device__ __forceinline__ void tmem_ld16_nc(uint32_t taddr, float (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]),
          "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait_dep(float (&v)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]),
                   "+f"(v[8]), "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15]));
}
// orders the uses of a second register block after a wait that named another block
__device__ __forceinline__ void tmem_dep(float (&v)[16]) {
    asm volatile("" : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]),
                      "+f"(v[8]), "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15]));
}
__device__ __forceinline__ void tmem_st16_nc(uint32_t taddr, const float (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]), "f"(v[8]),
        "f"(v[9]), "f"(v[10]), "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]));
}

// store 16 consecutive 32-bit columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
        "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])),
        "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])),
        "r"(__float_as_uint(v[11])), "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
        "r"(__float_as_uint(v[15]))
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- descriptors -------------------------------------------------------------------------------------------
constexpr uint64_t LAYOUT_SW128 = 2;        // UMMA::LayoutType::SWIZZLE_128B          (16-byte granule, 8-line atom)
constexpr uint64_t LAYOUT_SW128_B32 = 1;    // UMMA::LayoutType::SWIZZLE_128B_BASE32B  (32-byte granule, 4-line atom):
                                            // the only swizzle tcgen05 accepts for MN-major operands of 32-bit types

// smem operand descriptor: 128-byte-swizzled canonical layout.
//   K-major : rows (M or N index) are 128-byte lines of 32 tf32 along K; 8 rows = one 1024-byte atom; `sbo` = byte
//             stride between 8-row groups; `lbo` is ignored by the hardware for swizzled K-major (kept at 16).
//   MN-major: lines are 128 bytes = 32 consecutive M|N elements of one k; 8 k's = one atom; `lbo` = byte stride
//             between atoms along M|N, `sbo` = byte stride between atoms along K.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint64_t layout = LAYOUT_SW128) {
    uint64_t d = 0;
    d |= uint64_t((smem_addr >> 4) & 0x3FFF);
    d |= uint64_t((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= uint64_t((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= uint64_t(1) << 46;            // descriptor version (Blackwell)
    d |= layout << 61;
    return d;
}

// instruction descriptor for kind::tf32, fp32 accumulate
__host__ __device__ constexpr uint32_t make_idesc_tf32(int m, int n, bool a_mn_major, bool b_mn_major) {
    return (1u << 4)                      // c_format = F32
           | (2u << 7)                    // a_format = TF32
           | (2u << 10)                   // b_format = TF32
           | (uint32_t(a_mn_major) << 15) // a_major
           | (uint32_t(b_mn_major) << 16) // b_major
           | (uint32_t(n >> 3) << 17)     // n_dim
           | (uint32_t(m >> 4) << 24);    // m_dim
}

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    const uint32_t acc = accumulate ? 1u : 0u;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
// Warp-collective forms: the WHOLE warp executes the call (operands stay warp-uniform, so they can live in uniform
// registers instead of being converted per instruction); one elected lane issues.
__device__ __forceinline__ void umma_tf32_warp(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    const uint32_t acc = accumulate ? 1u : 0u;
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
// A operand from tensor memory (lane = M row, one 32-bit column per K element), B from shared memory
__device__ __forceinline__ void umma_tf32_ts_warp(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    const uint32_t acc = accumulate ? 1u : 0u;
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void umma_commit_warp(uint64_t* bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
        ::"r"(smem_u32(bar))
        : "memory");
}

// make the mbarrier observe completion of all previously issued tcgen05.mma of this thread
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- fp32 -> (tf32 hi, tf32-representable lo) split -----------------------------------------------------------
// hi is x rounded to tf32 precision (10 mantissa bits, low 13 bits zero); lo = x - hi is exact in fp32 and its own
// top bits carry the next ~11 bits of x.  hi*hi + hi*lo + lo*hi reproduces the fp32 product to ~2^-21.
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    hi = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);   // round to 11 significant bits
    lo = x - hi;                                                          // exact
}

// byte offset of element (row, col) inside a [rows][32 floats] tile stored as 128-byte lines with the 128B swizzle
// (16-byte chunk index XOR (row & 7)); the tile base must be 1024-byte aligned.
__device__ __forceinline__ uint32_t sw128_offset(uint32_t row, uint32_t col) {
    return row * 128u + ((((col >> 2) ^ (row & 7u)) & 7u) << 4) + ((col & 3u) << 2);
}

// MN-major fp32/tf32 operand (SWIZZLE_128B_BASE32B): element (mn, k) of an operand whose atoms (32 mn x 4 k, 512 B)
// are `lbo` bytes apart along M|N and `sbo` bytes apart along K.  Lines are 128 B (32 consecutive mn of one k); the
// 32-byte granule index is XORed with the line index inside the atom.  Atom bases must be 512-byte aligned.
__device__ __forceinline__ uint32_t mn32_offset(uint32_t mn, uint32_t k, uint32_t lbo, uint32_t sbo) {
    const uint32_t line = k & 3u;
    return (k >> 2) * sbo + (mn >> 5) * lbo + line * 128u + (((((mn & 31u) >> 3) ^ line) & 3u) << 5) + ((mn & 7u) << 2);
}

}  // namespace wxtc
