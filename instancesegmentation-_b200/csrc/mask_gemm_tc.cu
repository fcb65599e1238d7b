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

// ---------------------------------------------------------------------------------------------------------------
// Batched, pipelined variant: sigmoid(proto[b] @ coef[b]^T) for a whole batch in ONE persistent launch.
//
//   * a CTA owns a contiguous range of (image, 128-pixel tile) work items, so the image's coefficient operand (3 exact
//     tf32 parts, K-major SWIZZLE_128B) is built once and stays resident in shared memory while the CTA walks the tiles;
//   * the raw fp32 prototype tile (128 pixels x 32 channels = 16 KB, contiguous in [B, Ph*Pw, 32]) arrives by bulk TMA
//     (cp.async.bulk.shared::cluster.global + mbarrier complete_tx, SASS UBLKCP) into a two-deep ring: tile i + 1 is in
//     flight while tile i is split, multiplied and written out;
//   * 256 threads split the tile into the three tf32 parts (two threads per pixel row), ONE thread issues the 24
//     tcgen05.mma (M = 128, N = round16(m), K = 8; six exact partial products x four k-steps) into tensor memory and
//     commits to an mbarrier;
//   * epilogue: tcgen05.ld (warp w reads TMEM lanes 32 (w mod 4) .., column half w / 4) -> optional sigmoid -> the tile's
//     output image [rows, m] fp32 in shared memory -> ONE bulk asynchronous store of rows * m * 4 contiguous bytes
//     (the tile's rows are adjacent in [B, Ph*Pw, m]); the store is only waited for before the staging is overwritten.
//   mode kTcFootprint instead scatters sigmoid values into the packed per-detection footprint maps the compose kernel
//   reads (K3a's output layout), so the tensor-core product can stand in for sigmoid_map_kernel on the fused path
//   (ISB_TC_SMAP=1; measured in DESIGN.md section 4 -- it is not the default because the dense product computes 16 x
//   more elements than the footprints need).
// ---------------------------------------------------------------------------------------------------------------
constexpr int kTbThreads = 256;
enum { kTcLogits = 0, kTcSigmoid = 1, kTcFootprint = 2 };

struct TcBatchParams {
  const float* proto;      // [B, npix, 32]
  const float* coef;       // [B, coef_rows, 32]
  int64_t coef_rows;       // rows between images in coef
  const int32_t* counts;   // [B] detections per image (device) or nullptr -> m_fixed
  int32_t m_fixed;
  int32_t batch, npix, proto_w;
  int32_t mode;
  float* out;              // modes 0/1: [B, npix, m_fixed] dense
  const DetRec* rec;       // mode 2: [B, rec_stride] records (footprint origin / size / offset)
  int32_t rec_stride;
  float* smap;             // mode 2: [B, s_cap] packed footprint maps
  int64_t s_cap;
};

struct TbSmem {
  static constexpr size_t kRaw = (size_t)kTcM * 128;          // one raw fp32 tile
  static constexpr size_t kA = (size_t)kTcM * 128;            // one A part
  static constexpr size_t kB = (size_t)kTcNMax * 128;         // one B part
  static constexpr size_t off_raw = 0;
  static constexpr size_t off_a = 2 * kRaw;
  static constexpr size_t off_b = off_a + 3 * kA;
  static constexpr size_t off_out = off_b + 3 * kB;
  static constexpr size_t total = off_out + (size_t)kTcM * kTcNMax * 4 + 1024;  // + alignment slack
};

__global__ void __launch_bounds__(kTbThreads, 1) mask_logits_batch_tc_kernel(TcBatchParams p) {
  extern __shared__ unsigned char smem_dyn[];
  __shared__ __align__(8) uint64_t raw_bar[2];
  __shared__ __align__(8) uint64_t mma_bar;
  __shared__ uint32_t s_tmem;
  __shared__ DetRec s_rec[kTcNMax];
  unsigned char* base = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* sRaw = base + TbSmem::off_raw;
  unsigned char* sA = base + TbSmem::off_a;
  unsigned char* sB = base + TbSmem::off_b;
  float* sOut = reinterpret_cast<float*>(base + TbSmem::off_out);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tiles_per_img = (p.npix + kTcM - 1) / kTcM;
  const long long total_tiles = (long long)p.batch * tiles_per_img;
  const long long per_cta = (total_tiles + gridDim.x - 1) / gridDim.x;
  const long long t_begin = (long long)blockIdx.x * per_cta;
  const long long t_end = t_begin + per_cta < total_tiles ? t_begin + per_cta : total_tiles;
  if (t_begin >= t_end) return;  // (whole CTA: before any barrier or TMEM allocation)

  if (warp == 0) tc::tmem_alloc(&s_tmem, kTcCols);
  if (tid == 0) {
    mbar_init(&raw_bar[0], 1);
    mbar_init(&raw_bar[1], 1);
    mbar_init(&mma_bar, 1);
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = s_tmem;

  auto issue_raw = [&](long long tile, int buf) {  // thread 0 only
    const int b = (int)(tile / tiles_per_img), t = (int)(tile - (long long)b * tiles_per_img);
    const int rows = min(kTcM, p.npix - t * kTcM);
    const float* src = p.proto + ((size_t)b * p.npix + (size_t)t * kTcM) * ISB_NUM_COEF;
    mbar_expect_tx(&raw_bar[buf], (uint32_t)rows * 128u);
    bulk_g2s(sRaw + (size_t)buf * TbSmem::kRaw, src, (uint32_t)rows * 128u, &raw_bar[buf]);
  };
  if (tid == 0) issue_raw(t_begin, 0);

  uint32_t raw_phase[2] = {0u, 0u}, mma_phase = 0u;
  int cur_img = -1, m = 0, npad = 16;
  int it = 0;
  for (long long tile = t_begin; tile < t_end; ++tile, ++it) {
    const int buf = it & 1;
    const int b = (int)(tile / tiles_per_img), t = (int)(tile - (long long)b * tiles_per_img);
    const int row0 = t * kTcM, rows = min(kTcM, p.npix - row0);
    // the other raw buffer was consumed (converted) in the previous iteration: refill it with the next tile
    if (tid == 0 && tile + 1 < t_end) issue_raw(tile + 1, buf ^ 1);
    if (b != cur_img) {
      // new image: its coefficient operand (and, for the footprint mode, its detection records).  The previous tile's
      // MMAs have completed (waited for below), so sB is free.
      cur_img = b;
      m = p.counts ? p.counts[b] : p.m_fixed;
      m = max(0, min(m, kTcNMax));
      npad = max(16, (m + 15) & ~15);
      const float* cimg = p.coef + (size_t)b * p.coef_rows * ISB_NUM_COEF;
      for (int i = tid; i < npad * 8; i += kTbThreads) {
        const int n = i >> 3, c = i & 7;
        const float4 v = n < m ? __ldg(reinterpret_cast<const float4*>(cimg + (size_t)n * ISB_NUM_COEF) + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 p1, p2, p3;
        tc::split3_tf32(v.x, p1.x, p2.x, p3.x);
        tc::split3_tf32(v.y, p1.y, p2.y, p3.y);
        tc::split3_tf32(v.z, p1.z, p2.z, p3.z);
        tc::split3_tf32(v.w, p1.w, p2.w, p3.w);
        const uint32_t o = tc::sw128_offset(n, c);
        *reinterpret_cast<float4*>(sB + o) = p1;
        *reinterpret_cast<float4*>(sB + TbSmem::kB + o) = p2;
        *reinterpret_cast<float4*>(sB + 2 * TbSmem::kB + o) = p3;
      }
      if (p.mode == kTcFootprint)
        for (int j = tid; j < m; j += kTbThreads) s_rec[j] = p.rec[(size_t)b * p.rec_stride + j];
    }
    // ---- raw tile -> three tf32 parts, K-major SWIZZLE_128B (two threads per pixel row) -----------------------------
    mbar_wait(&raw_bar[buf], raw_phase[buf]);
    raw_phase[buf] ^= 1u;
    {
      const int r = tid >> 1, c0 = (tid & 1) * 4;
      const float4* src = reinterpret_cast<const float4*>(sRaw + (size_t)buf * TbSmem::kRaw + (size_t)r * 128);
#pragma unroll
      for (int c = c0; c < c0 + 4; ++c) {
        const float4 v = r < rows ? src[c] : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 p1, p2, p3;
        tc::split3_tf32(v.x, p1.x, p2.x, p3.x);
        tc::split3_tf32(v.y, p1.y, p2.y, p3.y);
        tc::split3_tf32(v.z, p1.z, p2.z, p3.z);
        tc::split3_tf32(v.w, p1.w, p2.w, p3.w);
        const uint32_t o = tc::sw128_offset(r, c);
        *reinterpret_cast<float4*>(sA + o) = p1;
        *reinterpret_cast<float4*>(sA + TbSmem::kA + o) = p2;
        *reinterpret_cast<float4*>(sA + 2 * TbSmem::kA + o) = p3;
      }
    }
    fence_proxy_async();  // generic-proxy smem writes (operands; raw reads done) -> async proxy (tensor core, next TMA refill)
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    if (m > 0) {
      if (tid == 0) {
        const uint32_t idesc = tc::idesc_tf32(kTcM, npad);
        const int pa[6] = {2, 0, 1, 1, 0, 0};
        const int pb[6] = {0, 2, 1, 0, 1, 0};
        bool acc = false;
#pragma unroll
        for (int q = 0; q < 6; ++q) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            tc::mma_tf32(tmem, tc::smem_desc_k128(sA + pa[q] * TbSmem::kA, k * 32), tc::smem_desc_k128(sB + pb[q] * TbSmem::kB, k * 32),
                         idesc, acc);
            acc = true;
          }
        }
        tc::mma_commit(&mma_bar);
      }
      mbar_wait(&mma_bar, mma_phase);
      mma_phase ^= 1u;
      tc::fence_after_sync();
      // ---- epilogue ---------------------------------------------------------------------------------------------
      const int q4 = warp & 3, half = warp >> 2;            // TMEM lane quarter this warp may read; column half it takes
      const int r = q4 * 32 + lane;
      const int c_lo = half ? ((npad / 2 + 7) & ~7) : 0, c_hi = half ? npad : ((npad / 2 + 7) & ~7);
      const uint32_t taddr = tmem + ((uint32_t)(q4 * 32) << 16);
      if (p.mode != kTcFootprint) {
        if (tid == 0) bulk_wait_read0();  // the previous tile's bulk store has finished reading sOut
        __syncthreads();
        for (int c = c_lo; c < c_hi; c += 8) {
          float v[8];
          tc::tmem_ld_32x32b_x8(taddr + c, v);
          if (p.mode == kTcSigmoid) {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = sigmoidf_ref(v[i]);
          }
          if (c + 8 <= m && (m & 3) == 0) {
            *reinterpret_cast<float4*>(sOut + (size_t)r * m + c) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(sOut + (size_t)r * m + c + 4) = make_float4(v[4], v[5], v[6], v[7]);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              if (c + i < m) sOut[(size_t)r * m + c + i] = v[i];
          }
        }
        tc::fence_before_sync();
        fence_proxy_async();
        __syncthreads();
        float* dst = p.out + ((size_t)b * p.npix + row0) * (size_t)m;
        const size_t bytes = (size_t)rows * m * 4;
        if (((reinterpret_cast<uintptr_t>(dst) | bytes) & 15) == 0) {
          if (tid == 0) {
            bulk_s2g(dst, sOut, (uint32_t)bytes);
            bulk_commit();
          }
        } else {
          for (int i = tid; i < rows * m; i += kTbThreads) dst[i] = sOut[i];
          __syncthreads();
        }
      } else {
        // scatter sigmoid(logit) of (pixel r, detection j) into detection j's packed footprint map if the pixel lies in it
        const int pix = row0 + r;
        const int py = pix / p.proto_w, px = pix - py * p.proto_w;
        float* S = p.smap + (size_t)b * p.s_cap;
        for (int c = c_lo; c < c_hi; c += 8) {
          float v[8];
          tc::tmem_ld_32x32b_x8(taddr + c, v);
          if (r < rows) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (c + i < m) {
                const DetRec d = s_rec[c + i];
                const int fx = px - d.pa, fy = py - d.pya;
                if (fx >= 0 && fx < d.fw && fy >= 0 && fy < d.fh) S[d.soff + fy * d.fw + fx] = sigmoidf_ref(v[i]);
              }
            }
          }
        }
        tc::fence_before_sync();
        __syncthreads();
      }
    }
  }
  if (tid == 0) bulk_wait0();
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, kTcCols);
}

int launch_mask_logits_batch(isb_handle* h, const TcBatchArgs& a, cudaStream_t s) {
  static_assert(kTcBatchMaxDet == kTcNMax, "one N tile per image");
  TcBatchParams p;
  p.proto = a.proto; p.coef = a.coef; p.coef_rows = a.coef_rows; p.counts = a.counts; p.m_fixed = a.m_fixed;
  p.batch = a.batch; p.npix = a.npix; p.proto_w = a.proto_w; p.mode = a.mode; p.out = a.out; p.rec = a.rec;
  p.rec_stride = a.rec_stride; p.smap = a.smap; p.s_cap = a.s_cap;
  const long long tiles = (long long)p.batch * ((p.npix + kTcM - 1) / kTcM);
  if (tiles <= 0) return ISB_OK;
  long long grid = h->num_sms;
  if (grid > tiles) grid = tiles;
  mask_logits_batch_tc_kernel<<<(unsigned)grid, kTbThreads, TbSmem::total, s>>>(p);
  ISB_LAUNCH_CHECK(h, "mask_logits_batch_tc_kernel");
  return ISB_OK;
}

int configure_mask_gemm_tc(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(mask_logits_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TcSmem::total));
  ISB_CUDA(h, cudaFuncSetAttribute(mask_logits_batch_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TbSmem::total));
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
