// Mask-logit GEMM on the 5th-generation tensor cores (tcgen05 + TMEM), utils_bbox.py:223 inner product:
//     logits[pix, j] = sum_k proto[pix, k] * coef[j, k]        (K = 32, fp32 in / fp32 out)
//
// The parity bar for mask logits is 1e-5, which a single kind::tf32 pass (10-bit mantissa, ~5e-4) cannot
// meet.  Each fp32 operand is therefore split EXACTLY into three tf32-representable parts
// (x = x1 + x2 + x3, 11 + 11 + 2 significant bits) and six partial products are accumulated in the
// fp32 accumulator in tensor memory:  a1b1 + a1b2 + a2b1 + a2b2 + a1b3 + a3b1  (dropped terms < 2^-33).
// Every partial product of two <= 11-bit parts is exact, so the result equals an fp32 dot product
// up to the accumulation rounding.
//
// One CTA = one 128-pixel x (<=128)-detection tile:  global -> registers -> split -> shared memory in the
// canonical K-major SWIZZLE_128B layout (rows of 32 fp32 = 128 B) -> 24 tcgen05.mma (M=128, N, K=8)
// issued by one thread -> tcgen05.commit on an mbarrier -> tcgen05.ld (32 lanes x 8 columns per warp)
// -> shared-memory transpose -> coalesced global store.
#include "tcgen05.cuh"

namespace isb {

constexpr int kTcThreads = 128;
constexpr int kTcM = 128;
constexpr int kTcNMax = 128;
constexpr int kTcCols = 128;  // TMEM columns allocated (power of two >= 32)

struct TcSmem {
  static constexpr size_t kA = (size_t)kTcM * 128;      // one A part
  static constexpr size_t kB = (size_t)kTcNMax * 128;   // one B part
  static constexpr size_t off_a = 0;
  static constexpr size_t off_b = 3 * kA;
  static constexpr size_t off_out = off_b + 3 * kB;
  static constexpr size_t total = off_out + (size_t)kTcM * (kTcNMax + 1) * 4 + 1024;  // + alignment slack
};

__global__ void __launch_bounds__(kTcThreads, 1)
mask_logits_tc_kernel(const float* __restrict__ proto, const float* __restrict__ coef, int64_t npix, int m,
                      float* __restrict__ out) {
  extern __shared__ unsigned char smem_dyn[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t s_tmem;
  unsigned char* base = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* sA = base + TcSmem::off_a;
  unsigned char* sB = base + TcSmem::off_b;
  float* sOut = reinterpret_cast<float*>(base + TcSmem::off_out);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t row0 = (int64_t)blockIdx.x * kTcM;
  const int n0 = blockIdx.y * kTcNMax;
  const int nvalid = min(kTcNMax, m - n0);
  const int npad = (nvalid + 15) & ~15;  // UMMA N: multiple of 16 for M = 128

  if (warp == 0) tc::tmem_alloc(&s_tmem, kTcCols);
  if (tid == 0) mbar_init(&bar, 1);

  // ---- stage A (128 prototype pixels) and B (npad detections), split into 3 tf32 parts ------------
  {
    const int64_t r = row0 + tid;
    const float4* src = reinterpret_cast<const float4*>(proto + r * ISB_NUM_COEF);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float4 v = r < npix ? __ldg(src + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 p1, p2, p3;
      tc::split3_tf32(v.x, p1.x, p2.x, p3.x);
      tc::split3_tf32(v.y, p1.y, p2.y, p3.y);
      tc::split3_tf32(v.z, p1.z, p2.z, p3.z);
      tc::split3_tf32(v.w, p1.w, p2.w, p3.w);
      const uint32_t o = tc::sw128_offset(tid, c);
      *reinterpret_cast<float4*>(sA + o) = p1;
      *reinterpret_cast<float4*>(sA + TcSmem::kA + o) = p2;
      *reinterpret_cast<float4*>(sA + 2 * TcSmem::kA + o) = p3;
    }
    for (int n = tid; n < npad; n += kTcThreads) {
      const float4* cs = reinterpret_cast<const float4*>(coef + (size_t)(n0 + n) * ISB_NUM_COEF);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float4 v = n < nvalid ? __ldg(cs + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 p1, p2, p3;
        tc::split3_tf32(v.x, p1.x, p2.x, p3.x);
        tc::split3_tf32(v.y, p1.y, p2.y, p3.y);
        tc::split3_tf32(v.z, p1.z, p2.z, p3.z);
        tc::split3_tf32(v.w, p1.w, p2.w, p3.w);
        const uint32_t o = tc::sw128_offset(n, c);
        *reinterpret_cast<float4*>(sB + o) = p1;
        *reinterpret_cast<float4*>(sB + TcSmem::kB + o) = p2;
        *reinterpret_cast<float4*>(sB + 2 * TcSmem::kB + o) = p3;
      }
    }
  }
  fence_proxy_async();      // generic-proxy smem writes -> visible to the tensor core (async proxy)
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = s_tmem;

  // ---- MMA: one thread issues 4 k-steps x 6 partial products -------------------------------------
  if (tid == 0) {
    const uint32_t idesc = tc::idesc_tf32(kTcM, npad);
    // (a part, b part) pairs, small terms first
    const int pa[6] = {2, 0, 1, 1, 0, 0};
    const int pb[6] = {0, 2, 1, 0, 1, 0};
    bool acc = false;
#pragma unroll
    for (int t = 0; t < 6; ++t) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {  // K = 8 tf32 = 32 bytes per step inside the 128-byte swizzled row
        const uint64_t da = tc::smem_desc_k128(sA + pa[t] * TcSmem::kA, k * 32);
        const uint64_t db = tc::smem_desc_k128(sB + pb[t] * TcSmem::kB, k * 32);
        tc::mma_tf32(tmem, da, db, idesc, acc);
        acc = true;
      }
    }
    tc::mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc::fence_after_sync();

  // ---- epilogue: TMEM -> registers -> shared (transpose) -> coalesced global ---------------------
  {
    const int r = warp * 32 + lane;
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < npad; c += 8) {
      float v[8];
      tc::tmem_ld_32x32b_x8(taddr + c, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) sOut[r * (kTcNMax + 1) + c + i] = v[i];
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  const int rows = (int)min((int64_t)kTcM, npix - row0);
  for (int idx = tid; idx < rows * nvalid; idx += kTcThreads) {
    const int r = idx / nvalid, n = idx - r * nvalid;
    out[(row0 + r) * m + n0 + n] = sOut[r * (kTcNMax + 1) + n];
  }
  if (warp == 0) tc::tmem_dealloc(tmem, kTcCols);
}

int configure_mask_gemm_tc(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(mask_logits_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TcSmem::total));
  return ISB_OK;
}

int launch_mask_logits_tc(isb_handle* h, const float* proto, const float* coef, int64_t npix, int m, float* out,
                          cudaStream_t s) {
  dim3 grid((unsigned)((npix + kTcM - 1) / kTcM), (unsigned)((m + kTcNMax - 1) / kTcNMax));
  mask_logits_tc_kernel<<<grid, kTcThreads, TcSmem::total, s>>>(proto, coef, npix, m, out);
  ISB_LAUNCH_CHECK(h, "mask_logits_tc_kernel");
  return ISB_OK;
}

}  // namespace isb
