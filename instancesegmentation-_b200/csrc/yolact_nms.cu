// K2 -- per-class top-k, Fast NMS, global top-N, box rescale.  One CTA per image.
//
// Replaces utils_bbox.py:55-99 (fast_non_max_suppression), :208-212 (second threshold) and
// :142-152 (yolact_correct_boxes).
//
//   1. sort the image's candidate keys (class asc, score desc, row asc)  -> the per-class sorted
//      lists of :61, restricted to the top_k window by rank (:63-64);
//   2. decode the candidates' boxes (same arithmetic as decode_boxes, :17-24);
//   3. Fast NMS: candidate j of a class is kept iff every IoU with a higher-ranked candidate of
//      the same class is <= nms_iou -- the column max of the upper-triangular IoU matrix
//      (:75-86) without ever materialising the matrix.  IoU uses the reference's fp32 op order
//      so keep sets are bit-exact; a NaN IoU suppresses (torch.max propagates NaN, NaN <= t is
//      false);
//   4. re-key the kept candidates (score desc, class asc, row asc), sort, keep max_det (:94-99);
//   5. emit boxes / scores / class ids / coefficients; in the fused path also scale+clamp the
//      boxes to pixels and precompute the integer crop ranges the mask kernel uses.
//
// Up to kSmemCand candidates are handled entirely in shared memory (keys 32 KB, boxes 64 KB,
// kept keys 32 KB); larger images fall back to global scratch with the same code (generic
// pointers).  Latency-bound integer/compare work: no tensor cores, no HBM traffic to speak of.
#include "isb_common.cuh"

namespace isb {

constexpr int kNmsThreads = 1024;
constexpr int kSmemCand = 4096;

__device__ void bitonic_sort_u64(uint64_t* k, int n /* power of two */) {
  for (int size = 2; size <= n; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int t = threadIdx.x; t < (n >> 1); t += blockDim.x) {
        int lo = 2 * t - (t & (stride - 1));
        int hi = lo + stride;
        bool asc = (lo & size) == 0;
        uint64_t a = k[lo], b = k[hi];
        if ((a > b) == asc) {
          k[lo] = b;
          k[hi] = a;
        }
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 load_box(const NmsParams& p, int b, uint32_t row) {
  if (p.boxes) return p.boxes[(size_t)b * p.rows + row];
  float4 pr = p.pred_box[(size_t)b * p.rows + row];
  float4 an = p.anchors[row];
  return decode_box(pr, an, 0.1f, 0.2f);
}

__global__ void __launch_bounds__(kNmsThreads, 1)
fast_nms_kernel(NmsParams p, const int32_t* __restrict__ cand_count, uint64_t* __restrict__ g_keys,
                uint64_t* __restrict__ g_det, float4* __restrict__ g_box, int64_t cap, int64_t stride,
                float* __restrict__ det_coef_ws, DetBounds* __restrict__ det_bounds_ws, int ws_det_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_kept;

  const int b = blockIdx.x;
  const int tid = threadIdx.x;
  const int64_t total = cand_count[b];
  if (total > cap) {  // candidate buffer overflowed: report, never guess
    if (tid == 0) p.out_count[b] = -1;
    return;
  }
  const int n = (int)total;
  if (n == 0) {
    if (tid == 0) p.out_count[b] = 0;
    return;
  }
  const int np2 = next_pow2(n);
  const bool in_smem = np2 <= kSmemCand;
  uint64_t* keys = in_smem ? reinterpret_cast<uint64_t*>(smem_raw) : g_keys + (size_t)b * stride;
  uint64_t* dkeys = in_smem ? reinterpret_cast<uint64_t*>(smem_raw) + kSmemCand : g_det + (size_t)b * stride;
  float4* cbox = in_smem ? reinterpret_cast<float4*>(smem_raw + 2 * kSmemCand * sizeof(uint64_t))
                         : g_box + (size_t)b * stride;

  // ---- 1. sort candidates -------------------------------------------------------------
  {
    const uint64_t* src = g_keys + (size_t)b * stride;
    if (in_smem) {
      for (int i = tid; i < np2; i += kNmsThreads) keys[i] = i < n ? src[i] : kPadKey;
    } else {
      for (int i = n + tid; i < np2; i += kNmsThreads) keys[i] = kPadKey;
    }
    if (tid == 0) s_kept = 0;
  }
  bitonic_sort_u64(keys, np2);

  // ---- 2. decode boxes of the candidates inside their class window ----------------------
  for (int i = tid; i < n; i += kNmsThreads) cbox[i] = load_box(p, b, cand_row(keys[i]));
  __syncthreads();

  // ---- 3. Fast NMS ----------------------------------------------------------------------
  const bool thr_ok = 0.0f <= p.nms_iou;  // column max of a triu'd matrix is >= 0 (:80-86)
  for (int j = tid; j < n; j += kNmsThreads) {
    const uint64_t kj = keys[j];
    const uint64_t cls_lo = (uint64_t)cand_class(kj) << (32 + kRowBits);
    // first index of this class: lower_bound over the sorted keys
    int lo = 0, hi = j;
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (keys[mid] < cls_lo) lo = mid + 1; else hi = mid;
    }
    const int start = lo;
    bool keep = thr_ok && (j - start) < p.top_k;
    if (keep) {
      const float4 bj = cbox[j];
      const float aj = box_area(bj);
      for (int i = start; i < j; ++i) {
        const float4 bi = cbox[i];
        float v = iou_pair(bi, box_area(bi), bj, aj);
        if (!(v <= p.nms_iou)) { keep = false; break; }
      }
    }
    if (keep) {
      int slot = atomicAdd(&s_kept, 1);
      dkeys[slot] = det_key_from_cand(kj);
    }
  }
  __syncthreads();

  // ---- 4. global order: score desc, class asc, row asc; keep max_det --------------------
  const int kept = s_kept;
  const int kp2 = next_pow2(kept);
  for (int i = kept + tid; i < kp2; i += kNmsThreads) dkeys[i] = kPadKey;
  bitonic_sort_u64(dkeys, kp2);
  const int m = min(kept, p.max_det);

  // ---- 5. outputs -----------------------------------------------------------------------
  const float wf = (float)p.out_w, hf = (float)p.out_h;
  for (int d = tid; d < m; d += kNmsThreads) {
    const uint64_t k = dkeys[d];
    const uint32_t row = det_row(k);
    float4 bx = load_box(p, b, row);
    if (p.correct) {
      bx = correct_box(bx, wf, hf);
      DetBounds db;
      crop_range(bx.x, bx.z, p.out_w, db.xlo, db.xhi);
      crop_range(bx.y, bx.w, p.out_h, db.ylo, db.yhi);
      det_bounds_ws[(size_t)b * ws_det_stride + d] = db;
    }
    reinterpret_cast<float4*>(p.out_box)[(size_t)b * p.max_det + d] = bx;
    p.out_score[(size_t)b * p.max_det + d] = det_score(k);
    p.out_class[(size_t)b * p.max_det + d] = (int64_t)det_class(k);
  }
  // coefficient rows: one warp-quarter (8 lanes x float4) per detection
  {
    float* dst_base = p.out_coef ? p.out_coef + (size_t)b * p.max_det * ISB_NUM_COEF
                                 : det_coef_ws + (size_t)b * ws_det_stride * ISB_NUM_COEF;
    for (int t = tid; t < m * 8; t += kNmsThreads) {
      int d = t >> 3, q = t & 7;
      const uint32_t row = det_row(dkeys[d]);
      const float4* src = reinterpret_cast<const float4*>(p.coef + ((size_t)b * p.rows + row) * ISB_NUM_COEF);
      reinterpret_cast<float4*>(dst_base + (size_t)d * ISB_NUM_COEF)[q] = src[q];
    }
  }
  if (tid == 0) p.out_count[b] = m;
}

static constexpr size_t kNmsSmem = (size_t)kSmemCand * (8 + 8 + 16);

int configure_fast_nms(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(fast_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kNmsSmem));
  return ISB_OK;
}

int launch_fast_nms(isb_handle* h, const NmsParams& p, cudaStream_t s) {
  const size_t b0 = (size_t)h->ws_b0, cap2 = (size_t)h->cand_cap_pow2;
  fast_nms_kernel<<<p.batch, kNmsThreads, kNmsSmem, s>>>(
      p, h->cand_count + b0, h->cand_keys + b0 * cap2, h->det_keys + b0 * cap2, h->cand_box + b0 * cap2,
      h->lim.cand_capacity, h->cand_cap_pow2, h->det_coef + b0 * h->lim.max_det * ISB_NUM_COEF,
      h->det_bounds + b0 * h->lim.max_det, h->lim.max_det);
  ISB_LAUNCH_CHECK(h, "fast_nms_kernel");
  return ISB_OK;
}

}  // namespace isb
