// sweep_common.cuh — mbarrier / bulk-copy (TMA 1-D) primitives and address helpers shared by the row-ring sweep kernels
// (backward_sweep.cu: DFMA evaluation, both arithmetic forms; backward_sweep_mma.cu: continuation values on FP64 mma.sync).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace lsmc {

// base + row * pitch_bytes as one IMAD.WIDE (the compiler otherwise widens the 32-bit pitch and adds a zero high word)
__device__ __forceinline__ const char *row_address(const char *base, uint32_t row, uint32_t pitch_bytes) {
  const char *p;
  asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(p) : "r"(row), "r"(pitch_bytes), "l"(base));
  return p;
}
// Bulk asynchronous copies (the TMA engine's 1-D form) complete on an mbarrier by byte count: one instruction moves a whole
// 1 KB row or a whole sub-tile of point records, no thread holds registers for it and nothing crosses the LSU data pipe.
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_load(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_prefetch_l2(const void *src_gmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}

}  // namespace lsmc
