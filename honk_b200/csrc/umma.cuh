// Thin inline-PTX wrappers for the sm_100a pieces the conv kernels use: tcgen05 (UMMA) with TMEM
// accumulators, mbarriers and TMA tensor loads.  No CUTLASS; encodings follow the PTX ISA
// (shared-memory matrix descriptor and kind::f16 instruction descriptor).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace honk {
namespace umma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// K-major, no-swizzle ("interleaved") operand: core matrix = 8 rows x 16 bytes stored as 128
// contiguous bytes; SBO = byte distance between 8-row groups (M/N direction); LBO = byte distance
// between the two 16-byte K chunks of one MMA (K = 16 half-precision elements).
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;   // descriptor version (Blackwell)
  return d;                              // base_offset 0, lbo_mode 0, layout SWIZZLE_NONE
}

// kind::f16 instruction descriptor: D = f32, A/B = f16 (fmt 0) or bf16 (fmt 1), both K-major.
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int ab_fmt) {
  return (1u << 4) | (static_cast<uint32_t>(ab_fmt) << 7) | (static_cast<uint32_t>(ab_fmt) << 10) |
         (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}

__device__ __forceinline__ void mma_f16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Arrives on the mbarrier once all previously issued MMAs of this thread have completed.
__device__ __forceinline__ void commit(uint32_t mbar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 32 lanes x 8 consecutive 32-bit columns: thread i of the warp gets lane (base + i).
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- mbarrier -------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t mbar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(mbar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t mbar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok)
               : "r"(mbar), "r"(parity)
               : "memory");
  return ok;
}
// Non-blocking probe of a phase (mbarrier.test_wait never suspends): used to look one task ahead.
__device__ __forceinline__ bool mbar_test_wait(uint32_t mbar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(mbar), "r"(parity)
      : "memory");
  return ok;
}
// Two probes whose latencies overlap (one result: both phases complete).
__device__ __forceinline__ bool mbar_test_wait2(uint32_t mbar_a, uint32_t parity_a, uint32_t mbar_b, uint32_t parity_b) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p, q;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 q, [%3], %4;\nand.pred p, p, q;\nselp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(mbar_a), "r"(parity_a), "r"(mbar_b), "r"(parity_b)
      : "memory");
  return ok;
}
// Three probes in flight at once (one result: all three phases complete).
__device__ __forceinline__ bool mbar_test_wait3(uint32_t mbar_a, uint32_t parity_a, uint32_t mbar_b, uint32_t parity_b,
                                                uint32_t mbar_c, uint32_t parity_c) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p, q, r;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 q, [%3], %4;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 r, [%5], %6;\n"
      "and.pred p, p, q;\nand.pred p, p, r;\nselp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(mbar_a), "r"(parity_a), "r"(mbar_b), "r"(parity_b), "r"(mbar_c), "r"(parity_c)
      : "memory");
  return ok;
}
// Bounded wait: a protocol bug (e.g. a TMA byte count that never arrives) traps instead of
// hanging the GPU.  try_wait suspends in hardware, so the poll count stays small in normal runs.
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
  for (uint32_t i = 0; !mbar_try_wait(mbar, parity); ++i)
    if (i > (1u << 24)) __trap();
}

// generic-proxy writes (st.shared) -> visible to the async proxy (UMMA / TMA reads)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMA tiled loads (coordinates innermost first) -------------------------------------------
__device__ __forceinline__ void tma_load_4d(uint32_t smem_dst, const void* tmap, uint32_t mbar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_dst),
      "l"(tmap), "r"(mbar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d(uint32_t smem_dst, const void* tmap, uint32_t mbar, int c0, int c1,
                                            int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_dst),
      "l"(tmap), "r"(mbar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

}  // namespace umma
}  // namespace honk
