// tc5.cuh — thin inline-PTX layer over the Blackwell tensor-core path (tcgen05.mma + TMEM + mbarrier), sm_100a only.
// Operand tiles use the no-swizzle ("interleaved") canonical shared-memory layout in a chunk-major arrangement:
//   a [ROWS x K] fp16 tile stores the 16-byte chunk (row r, k-chunk kc = k/8) at byte (kc*ROWS + r)*16.
// Read K-major (row = M or N index, K contiguous inside the chunk):  LBO = ROWS*16 (next k-chunk), SBO = 128 (next 8 rows).
// The same bytes read MN-major (the chunk's 8 elements run along M/N, rows run along K) describe the TRANSPOSED
// operand: LBO = 128 (next 8 k), SBO = ROWS*16 (next 8 m/n) — used for the weight-gradient products.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

namespace tc5 {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n"
        "W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra D;\n bra W;\n"
        "D:\n}" ::"r"(bar), "r"(parity) : "memory");
}

// ---- proxies / fences
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- TMEM allocation (one full warp calls these)
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// ---- descriptors
enum { MAJOR_K = 0, MAJOR_MN = 1 };
// shared-memory matrix descriptor, no swizzle (cute::UMMA::SmemDescriptor bit layout, version 1)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
// instruction descriptor for kind::f16: fp16 A/B, fp32 accumulate (cute::UMMA::InstrDescriptor bit layout)
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N, int a_major, int b_major) {
    return (1u << 4)                        // c_format = F32
           | (0u << 7) | (0u << 10)         // a_format = b_format = F16
           | ((uint32_t)a_major << 15) | ((uint32_t)b_major << 16) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]; one thread issues
__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
        " tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}
// arrive on the mbarrier when every MMA issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// ---- TMEM -> registers: the warp's 32 lanes x 16 consecutive fp32 columns (thread i gets lane 32*(warp%4)+i)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// ---- chunk-major tile addressing
__device__ __forceinline__ uint32_t chunk_off(int rows, int r, int kc) { return (uint32_t)(kc * rows + r) * 16u; }

__device__ __forceinline__ uint4 pack8(const float (&v)[16], int o) {
    __half2 a = __floats2half2_rn(v[o + 0], v[o + 1]), b = __floats2half2_rn(v[o + 2], v[o + 3]);
    __half2 c = __floats2half2_rn(v[o + 4], v[o + 5]), d = __floats2half2_rn(v[o + 6], v[o + 7]);
    uint4 u;
    u.x = *reinterpret_cast<uint32_t *>(&a);
    u.y = *reinterpret_cast<uint32_t *>(&b);
    u.z = *reinterpret_cast<uint32_t *>(&c);
    u.w = *reinterpret_cast<uint32_t *>(&d);
    return u;
}

// ---- extensions used by the warp-specialised field kernels (cfield5.cu, render5.cu) ---------------------------------
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// wait with a watchdog: a protocol bug traps after ~2 s instead of hanging the GPU.  The spin loop is out of line so that
// the many wait sites of the warp-specialised kernels cost two instructions each in the instruction cache.
static __device__ __noinline__ void mbar_spin_wd(uint32_t bar, uint32_t parity) {
    const long long t0 = clock64();
    while (!mbar_try(bar, parity)) {
        if (clock64() - t0 > 4000000000ll) {
            printf("tc5: mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x, (int)threadIdx.x, bar, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void mbar_wait_wd(uint32_t bar, uint32_t parity) {
    if (!mbar_try(bar, parity)) mbar_spin_wd(bar, parity);
}

// instruction descriptor for kind::tf32: tf32 A/B (fp32 containers, low 13 mantissa bits ignored), fp32 accumulate
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N, int a_major, int b_major) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_major << 15) | ((uint32_t)b_major << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[tmem] * B[smem]  (A: lane = row, column = k, one 32-bit element per column; 8 columns per instruction)
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
        " tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}
__device__ __forceinline__ void mma_tf32_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
        " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}

// registers -> TMEM: the warp's 32 lanes x 16 consecutive 32-bit columns (thread i writes lane 32*(warp%4)+i)
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st4(uint32_t taddr, const uint32_t (&r)[4]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3])
                 : "memory");
}
__device__ __forceinline__ void tmem_ld4_nowait(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// TMEM -> registers without the wait (issue several, then tmem_ld_wait())
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}

// fp32 -> (hi, lo) TF32 pair with hi + lo = x to ~2^-22 relative (both exactly representable in TF32).  Round-to-nearest
// (ties away) done with integer arithmetic: `cvt.rna.tf32.f32` compiles to 4 instructions per value on sm_100a (an
// inf/nan guard around the same add-and-mask); the operands here are finite.
__device__ __forceinline__ uint32_t rna_tf32(float x) { return (__float_as_uint(x) + 0x1000u) & 0xffffe000u; }
__device__ __forceinline__ void split_tf32(float x, uint32_t &hi, uint32_t &lo) {
    hi = rna_tf32(x);
    lo = rna_tf32(x - __uint_as_float(hi));
}

// TF32 operand tiles in shared memory use the same chunk-major no-swizzle layout: a [ROWS x K] fp32 tile stores the 16-byte
// chunk (row r, k-chunk kc = k/4) at byte (kc*ROWS + r)*16; one MMA (K = 8) covers two chunks.
__device__ __forceinline__ uint64_t tf32_kmajor_desc(uint32_t tile_addr, int rows, int kstep) {
    return smem_desc(tile_addr + (uint32_t)kstep * 2u * (uint32_t)rows * 16u, (uint32_t)rows * 16u, 128u);
}

}  // namespace tc5
