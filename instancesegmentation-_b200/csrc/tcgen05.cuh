// Minimal hand-written tcgen05 / TMEM helpers for sm_100a (inline PTX; SASS: UTCHMMA-family,
// LDTM, UTCBAR).  Only what the mask GEMM needs: single-CTA (cta_group::1) kind::tf32 MMA with both
// operands in shared memory (K-major, 128-byte swizzle), fp32 accumulation in tensor memory.
#pragma once

#include <stdint.h>

#include "isb_common.cuh"

namespace isb {
namespace tc {

// ---- tensor memory ------------------------------------------------------------------------
// ncols: power of two >= 32.  Executed by ONE full warp; the base address lands in *smem_slot.
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- descriptors --------------------------------------------------------------------------
// Shared-memory matrix descriptor, K-major, SWIZZLE_128B, rows of exactly 128 bytes (32 fp32):
//   bits [0,14)  start address >> 4        bits [16,30) leading byte offset >> 4 (unused for swizzled K-major: 1)
//   bits [32,46) stride byte offset >> 4 = 1024 B between 8-row groups      bits [46,48) version = 1 (Blackwell)
//   bits [61,64) layout type = 2 (SWIZZLE_128B)
// (cute/arch/mma_sm100_desc.hpp `SmemDescriptor`, canonical layout ((8,n),2):((8,SBO),1) in 16-byte units.)
__device__ __forceinline__ uint64_t smem_desc_k128(const void* tile_base_1024_aligned, uint32_t byte_offset_in_row) {
  const uint32_t addr = smem_u32(tile_base_1024_aligned) + byte_offset_in_row;
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Instruction descriptor for kind::tf32, fp32 accumulate, both operands K-major
// (cute/arch/mma_sm100_desc.hpp `InstrDescriptor`): c_format [4,6) = 1 (F32), a_format [7,10) = 2 (TF32),
// b_format [10,13) = 2, a/b major [15],[16] = 0 (K), n_dim [17,23) = N >> 3, m_dim [24,29) = M >> 4.
__host__ __device__ constexpr uint32_t idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]^T ; issued by ONE thread.
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}
// Arrive on an mbarrier once all previously issued MMAs of this thread have completed.
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- TMEM -> registers: this warp's 32 lanes x 8 consecutive fp32 columns -------------------
__device__ __forceinline__ void tmem_ld_32x32b_x8(uint32_t taddr, float v[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// ---- operand staging ------------------------------------------------------------------------
// Byte offset of (row, 16-byte chunk) inside a K-major SWIZZLE_128B tile with 128-byte rows:
// Swizzle<3,4,3>: address bits [4,7) ^= bits [7,10).
__device__ __forceinline__ uint32_t sw128_offset(int row, int chunk) { return (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4); }

// Exact three-way split of an fp32 into tf32-representable parts: x == x1 + x2 + x3, each with <= 11
// significant bits, so every partial product of two parts is exact in the tensor core's fp32 accumulate.
__device__ __forceinline__ void split3_tf32(float x, float& x1, float& x2, float& x3) {
  x1 = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
  const float r = x - x1;  // exact
  x2 = __uint_as_float(__float_as_uint(r) & 0xffffe000u);
  x3 = r - x2;             // exact, <= 2 significant bits
}

}  // namespace tc
}  // namespace isb
