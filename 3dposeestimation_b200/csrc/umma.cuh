/*
 * umma.cuh - thin inline-PTX layer over the sm_100a tensor-core path (tcgen05.mma, TMEM, mbarrier,
 * bulk async copies) used by the tensor kernels of this library.  Encodings follow the PTX ISA /
 * CUTLASS cute/arch/mma_sm100_desc.hpp bit layouts; nothing here is a library call.
 *
 * Shared-memory operand layout used throughout: K-major, no swizzle ("interleave").  In 16-byte
 * units an operand is ((8, m), 2) : ((1, SBO), LBO):
 *     - a core matrix is 8 rows x 16 bytes, rows 16 bytes apart (128 contiguous bytes);
 *     - SBO = byte distance between core matrices that are adjacent in the M/N direction;
 *     - LBO = byte distance between the two core matrices that make one K step (32 bytes of K).
 * The tensor core only computes addresses from these three numbers, so core matrices may overlap:
 * the convolution kernels exploit that to read im2col windows straight out of an HWC image
 * (LBO = one pixel, SBO = eight pixels) without materialising im2col.
 */
#pragma once
#include <cstdint>
#include <cuda_fp16.h>

namespace pe {
namespace umma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

/* ------------------------------------------------------------ descriptors --- */
/* shared-memory matrix descriptor: start [0,14) lbo [16,30) sbo [32,46) (all >>4), version=1 at [46,48),
 * layout type SWIZZLE_NONE = 0 at [61,64) */
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr >> 4) & 0x3FFFu);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    return d;
}
/* same descriptor, start address advanced by `bytes` (multiple of 16; no carry out of the 14-bit field
 * as long as the operand stays inside the 256 KB shared window) */
__device__ __forceinline__ uint64_t desc_advance(uint64_t d, uint32_t bytes) { return d + (bytes >> 4); }

enum { FMT_F16 = 0, FMT_BF16 = 1, FMT_TF32 = 2 };
/* instruction descriptor for kind::f16 / kind::tf32, fp32 accumulate, A and B both K-major */
__host__ __device__ constexpr uint32_t instr_desc(int m, int n, int a_fmt, int b_fmt) {
    return (1u << 4) | (static_cast<uint32_t>(a_fmt) << 7) | (static_cast<uint32_t>(b_fmt) << 10) |
           (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}

/* ------------------------------------------------------------------- TMEM --- */
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot) {   /* whole warp, .sync.aligned */
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {      /* whole warp, same warp as alloc */
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(COLS) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
/* generic-proxy shared-memory writes -> visible to the async proxy (tensor core / bulk copies) */
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

/* 32 lanes x 32 bit, N consecutive columns -> N registers per thread (thread = lane of its warp's quarter) */
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

/* -------------------------------------------------------------------- MMA --- */
/* one lane of a converged warp (elect.sync).  Issue tcgen05.mma / commit as
 *     if (elect_one()) { mma_f16(...); ... }  __syncwarp();
 * from code the WHOLE warp executes: the operands are then warp-uniform values (uniform registers) and ptxas emits a
 * plain predicated UTCHMMA.  Inside an `if (lane == 0)` region it instead wraps every MMA in an
 * ELECT / R2UR.BROADCAST / BRA.U.ANY loop that costs ~100 cycles per instruction. */
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

/* D[tmem] (+)= A[smem] * B[smem]; one thread issues for the CTA */
__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
/* kind::f8f6f4 with 8-bit operands (E4M3 = format code 0, the same instruction-descriptor bits as FMT_F16): K = 32 per
 * instruction, a core-matrix row is 16 bytes = 16 elements */
__device__ __forceinline__ void mma_f8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
/* all MMAs issued so far by this thread arrive (once) on the mbarrier when they have completed;
 * implies tcgen05.fence::before_thread_sync */
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

/* --------------------------------------------------------------- mbarrier --- */
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
/* bounded wait: a protocol bug becomes a reported error (false) instead of a hung GPU */
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, uint32_t max_spins = (1u << 20)) {
#pragma unroll 1
    for (uint32_t i = 0; i < max_spins; ++i)
        if (mbar_try_wait(bar, parity)) return true;
    return false;
}

/* the same for a role with slack (e.g. a loader that runs ahead): sleeps between polls so that its spinning
 * does not take issue slots from the warps on the critical path */
__device__ __forceinline__ bool mbar_wait_relaxed(uint64_t* bar, uint32_t parity, uint32_t ns = 256, uint32_t max_spins = (1u << 18)) {
#pragma unroll 1
    for (uint32_t i = 0; i < max_spins; ++i) {
        if (mbar_try_wait(bar, parity)) return true;
        __nanosleep(ns);
    }
    return false;
}

/* ------------------------------------------------------------ bulk copies --- */
/* 1-D bulk async copy global -> shared (TMA engine, no tensor map); bytes % 16 == 0, both 16-byte aligned */
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

/* ---------------------------------------------------------- fp16 splitting --- */
/* x ~= hi + lo with hi = fp16(x), lo = fp16(x - hi): 22 significant bits when both are normal */
__device__ __forceinline__ void split_f16(float x, __half& hi, __half& lo) {
    hi = __float2half_rn(x);
    lo = __float2half_rn(x - __half2float(hi));
}

}  // namespace umma
}  // namespace pe
