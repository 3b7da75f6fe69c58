// Inline-PTX helpers shared by the sm_100a kernels: mbarrier, bulk async copy (TMA), warp-level
// int8 MMA, ldmatrix / stmatrix, saturating packs.  Fragment layouts were checked on the box with
// tools/probe_ldsm.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mibmi {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier -------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done != 0;
}
// suspend-time hint of try_wait: a waiting warp sleeps in hardware (no issue slots) until the phase completes
// or the hint expires, instead of spinning through the scheduler
#ifndef MIBMI_WAIT_HINT_NS
#define MIBMI_WAIT_HINT_NS 20000u
#endif
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    // test_wait first: a phase that has already completed costs one probe instead of the (suspending)
    // try_wait round trip -- measured +3 % on the whole kernel
    if (mbar_test(bar, parity)) return;
    uint32_t done = 0;
    for (uint32_t spins = 0; !done; ++spins) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity), "r"(MIBMI_WAIT_HINT_NS) : "memory");
        if (spins > (1u << 22)) __trap();   // never hang the box: a lost arrival becomes a launch failure
    }
}
// same with a sleep between probes: for roles that wait most of the time (their polling would otherwise take
// issue slots from the working warps)
__device__ __forceinline__ void mbar_wait_sleep(uint32_t bar, uint32_t parity, uint32_t ns) {
    for (uint32_t spins = 0; !mbar_test(bar, parity); ++spins) {
        __nanosleep(ns);
        if (spins > (1u << 24)) __trap();
    }
}
// pure hardware wait: mbarrier.try_wait suspends the thread until the phase completes or the (implementation-defined) time
// limit expires -- no instruction is issued while it sleeps, the loop only re-arms after a time-out
__device__ __forceinline__ void mbar_wait_hw(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spins = 0; !done; ++spins) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity), "r"(MIBMI_WAIT_HINT_NS) : "memory");
        if (spins > (1u << 22)) __trap();
    }
}
// MODE 0: test_wait + nanosleep polling (round 1), 1: test_wait, then the suspending try_wait, 2: try_wait only
template <int MODE>
__device__ __forceinline__ void mbar_wait_mode(uint32_t bar, uint32_t parity, uint32_t ns) {
    if (MODE == 0) mbar_wait_sleep(bar, parity, ns);
    else if (MODE == 1) mbar_wait(bar, parity);
    else mbar_wait_hw(bar, parity);
}
// TMA bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// TMA 2-D tile load through a tensor map (SASS: UTMALDG), completion on an mbarrier
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(tmap), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void* tmap, int c0, int c1, int c2, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- tcgen05 (5th-generation tensor core, accumulators in TMEM) ----------------------------------
// descriptor / layout facts were verified on the box with tools/probe_umma.cu
__device__ __forceinline__ void tmem_alloc(uint32_t smem_result, uint32_t cols) {      // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {          // the same warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address, leading / stride byte offsets
// (all >> 4), version 1 (Blackwell), layout type (0 = no swizzle, 2 = 128-byte swizzle)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout_type) {
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32) | ((uint64_t)1 << 46) | ((uint64_t)layout_type << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor) of kind::i8: S32 accumulator, signed int8 A and B,
// A MN-major (the [channel][sample] tile), B K-major, N >> 3 at bit 17, M >> 4 at bit 24
__device__ __forceinline__ constexpr uint32_t umma_idesc_i8(int M, int N) {
    return (2u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem], issued by ONE thread (SASS: UTCIMMA)
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                 ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// same with the descriptors given as (low word, high word): the issuing thread only adds to the low word
// (start-address field, units of 16 bytes) between instructions
__device__ __forceinline__ void umma_i8_lohi(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                             uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %6, 0;\n\t"
                 "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], da, db, %5, {%7, %7, %7, %7}, p;\n\t}"
                 ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// one lane of a converged warp (the warp keeps warp-uniform control flow around the tcgen05 issue)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// arrive on an mbarrier when every tcgen05.mma issued so far by this thread has completed (SASS: UTCBAR)
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];" ::"r"(bar) : "memory");
}
// 16 TMEM lanes x 16 columns -> 8 registers in the m16n8 accumulator-fragment order (SASS: LDTM.16dp256bit.x2):
// r[4 nt + c] = (lane g + 8 (c >> 1), column 8 nt + 2 q + (c & 1))
__device__ __forceinline__ void tmem_ld_16x256b_x2(uint32_t (&r)[8], uint32_t taddr) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
}
// 16 TMEM lanes x 8 columns -> 4 registers: r[c] = (lane g + 8 (c >> 1), column 2 q + (c & 1))   (SASS: LDTM.16dp256bit)
__device__ __forceinline__ void tmem_ld_16x256b_x1(uint32_t (&r)[4], uint32_t taddr) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- warp-level int8 MMA (SASS: IMMA.16832.S8.S8 / IMMA.16816.S8.S8) --------------------------
//   A (row)  a0: (row g, k 4q..4q+3)  a1: (row g+8, k 4q..)  a2: (row g, k 16+4q..)  a3: (row g+8, k 16+4q..)
//   B (col)  b0: (k 4q..4q+3, col g)  b1: (k 16+4q.., col g)
//   C        c0, c1: (row g, cols 2q, 2q+1)   c2, c3: (row g+8, cols 2q, 2q+1)        g = lane>>2, q = lane&3
__device__ __forceinline__ void mma_k32(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_k16(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a0), "r"(a1), "r"(b0));
}

// ---- ldmatrix / stmatrix ----------------------------------------------------------------------
// four 8x16-byte matrices; lane l supplies the address of row (l & 7) of matrix (l >> 3); register j of
// thread (g, q) = bytes 4q..4q+3 of row g of matrix j
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
// two 8x16-byte matrices (lanes 0-7 / 8-15 supply the row addresses; the others must still be valid)
__device__ __forceinline__ void ldsm_x2(uint32_t (&r)[2], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];"
                 : "=r"(r[0]), "=r"(r[1]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void stsm_x4(uint32_t addr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
    asm volatile("stmatrix.sync.aligned.m8n8.x4.shared.b16 [%0], {%1,%2,%3,%4};"
                 ::"r"(addr), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
}
// transposing byte store of MMA accumulator fragments: four 16x8 byte matrices V; register j of thread (g, q)
// holds { V[g][2q], V[g][2q+1], V[g+8][2q], V[g+8][2q+1] } of matrix j (the m16n8 C-fragment order, saturated
// to bytes); memory receives V^T: row c (8 rows of 16 bytes, address from lane 8 j + c) = column c of V
__device__ __forceinline__ void stsm_t8_x4(uint32_t addr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
    asm volatile("stmatrix.sync.aligned.m16n8.x4.trans.shared.b8 [%0], {%1,%2,%3,%4};"
                 ::"r"(addr), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
}
__device__ __forceinline__ void stsm_t8_x2(uint32_t addr, uint32_t r0, uint32_t r1) {       // two matrices: addresses from lanes 0..15
    asm volatile("stmatrix.sync.aligned.m16n8.x2.trans.shared.b8 [%0], {%1,%2};" ::"r"(addr), "r"(r0), "r"(r1) : "memory");
}
// transposing byte load: 16 rows x 16 bytes per matrix, lane l (l < 16 per matrix; lanes 16..31 address the
// second matrix of .x2) supplies row l & 15.  Thread (g, q) receives per matrix M
//   r0 = { M[4q][g], M[4q+1][g], M[4q+2][g], M[4q+3][g] }      r1 = the same rows at column g + 8
__device__ __forceinline__ void ldsm_t_x1(uint32_t (&r)[2], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m16n16.x1.trans.shared.b8 {%0,%1}, [%2];"
                 : "=r"(r[0]), "=r"(r[1]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ldsm_t_x2(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}

// ---- saturating packs (SASS: I2IP.S8.S32.SAT) --------------------------------------------------
// bytes (v0, v1, v2, v3) -> one word, each saturated to [-128, 127]
__device__ __forceinline__ uint32_t pack4_sat(int v0, int v1, int v2, int v3) {
    uint32_t t, d;
    asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, 0;" : "=r"(t) : "r"(v3), "r"(v2));
    asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(v1), "r"(v0), "r"(t));
    return d;
}
// bytes (v0, v1) in the low half-word
__device__ __forceinline__ uint32_t pack2_sat(int v0, int v1) {
    uint32_t d;
    asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, 0;" : "=r"(d) : "r"(v1), "r"(v0));
    return d;
}

}  // namespace mibmi
