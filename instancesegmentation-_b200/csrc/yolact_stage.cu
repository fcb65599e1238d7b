// Small element-wise stage kernels behind the per-method entry points of BBoxUtility
// (decode_boxes, jaccard, yolact_correct_boxes, crop) plus the plain mask-logit product.
// All are coalesced float4 / one-thread-per-element passes: HBM-bound by construction.
#include "isb_common.cuh"

namespace isb {

// utils_bbox.py:11-25
__global__ void decode_boxes_kernel(const float4* __restrict__ pred, const float4* __restrict__ anchors, int64_t n,
                                    float var0, float var1, float4* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = decode_box(pred[i], anchors[i], var0, var1);
}

// utils_bbox.py:27-48 ; out[bt, i, j]
__global__ void jaccard_kernel(const float4* __restrict__ a, const float4* __restrict__ b, int batch, int na, int nb,
                               int iscrowd, float* __restrict__ out) {
  const int64_t total = (int64_t)batch * na * nb;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(t % nb);
    const int64_t r = t / nb;
    const int i = (int)(r % na);
    const int bt = (int)(r / na);
    const float4 ba = a[(size_t)bt * na + i], bb = b[(size_t)bt * nb + j];
    const float aa = box_area(ba), ab = box_area(bb);
    if (iscrowd) {
      float dx = fmaxf(fminf(ba.z, bb.z) - fmaxf(ba.x, bb.x), 0.0f);
      float dy = fmaxf(fminf(ba.w, bb.w) - fmaxf(ba.y, bb.y), 0.0f);
      out[t] = (dx * dy) / aa;
    } else {
      out[t] = iou_pair(ba, aa, bb, ab);
    }
  }
}

// utils_bbox.py:142-152
__global__ void correct_boxes_kernel(const float4* __restrict__ in, int64_t m, float wf, float hf,
                                     float4* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = correct_box(in[i], wf, hf);
}

// utils_bbox.py:154-168 ; masks [H,W,m] * inside(box_j)
__global__ void crop_kernel(const float* __restrict__ masks, const float4* __restrict__ boxes, int hgt, int wid, int m,
                            float* __restrict__ out) {
  const int64_t total = (int64_t)hgt * wid * m;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(t % m);
    const int64_t pix = t / m;
    const float col = (float)(pix % wid), row = (float)(pix / wid);
    const float4 bx = __ldg(boxes + j);
    const bool inside = (col >= bx.x) && (col < bx.z) && (row >= bx.y) && (row < bx.w);
    out[t] = masks[t] * (inside ? 1.0f : 0.0f);
  }
}

// pixel-space boxes -> per-detection records of the mask kernels: integer crop ranges, prototype footprint and
// the offset of the detection's sigmoid map (stage entry isb_mask_assemble; the fused path does this in the
// NMS kernel).  One CTA, m <= 1024.
__global__ void __launch_bounds__(1024) det_layout_kernel(const float4* __restrict__ boxes, int m, int hgt, int wid,
                                                          int proto_h, int proto_w, DetRec* __restrict__ out,
                                                          int32_t* __restrict__ total_out) {
  __shared__ int s_scan[32];
  const int j = threadIdx.x;
  DetRec rc;
  int items = 0;
  if (j < m) {
    const float4 bx = boxes[j];
    DetBounds db;
    crop_range(bx.x, bx.z, wid, db.xlo, db.xhi);
    crop_range(bx.y, bx.w, hgt, db.ylo, db.yhi);
    int pa, pya, fw, fh;
    det_footprint(db, (float)proto_h / (float)hgt, (float)proto_w / (float)wid, proto_h, proto_w, pa, pya, fw, fh);
    rc.xlo = db.xlo; rc.xhi = db.xhi; rc.ylo = db.ylo; rc.yhi = db.yhi;
    rc.pa = (int16_t)pa; rc.pya = (int16_t)pya; rc.fw = (int16_t)fw; rc.fh = (int16_t)fh;
    rc.j = j;
    items = fw * fh;
  }
  int total;
  rc.soff = cta_exclusive_scan(items, s_scan, &total);
  if (j < m) out[j] = rc;
  if (j == 0) *total_out = total;
}

// utils_bbox.py:104 torch.argmax(class_thre, dim=-1): first maximum of every row (a NaN entry wins, as in torch)
__global__ void row_argmax_kernel(const float* __restrict__ scores, int n, int ncols, int32_t* __restrict__ labels) {
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
    const float* row = scores + (size_t)r * ncols;
    float best = row[0];
    int arg = 0;
    for (int c = 1; c < ncols && best == best; ++c) {
      const float v = row[c];
      if (v > best || v != v) { best = v; arg = c; }
    }
    labels[r] = arg;
  }
}

// utils_bbox.py:126-139: rows of the kept candidates, already in final order
__global__ void trad_gather_kernel(const int64_t* __restrict__ keep, const int32_t* __restrict__ count, int max_det,
                                   const float4* __restrict__ boxes, const float* __restrict__ scores,
                                   const int32_t* __restrict__ labels, const float* __restrict__ coef,
                                   float4* __restrict__ out_box, float* __restrict__ out_score,
                                   int64_t* __restrict__ out_class, float* __restrict__ out_coef,
                                   int32_t* __restrict__ out_count) {
  const int m = min(*count, max_det);
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < m * 8; t += gridDim.x * blockDim.x) {
    const int d = t >> 3, q = t & 7;
    const int64_t i = keep[d];
    reinterpret_cast<float4*>(out_coef + (size_t)d * ISB_NUM_COEF)[q] = reinterpret_cast<const float4*>(coef + i * ISB_NUM_COEF)[q];
    if (q == 0) {
      out_box[d] = boxes[i];
      out_score[d] = scores[i];
      out_class[d] = (int64_t)labels[i];
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *out_count = m;
}

// utils_bbox.py:223 inner product: logits[pix, j] = sum_k proto[pix, k] * coef[j, k]
__global__ void mask_logits_kernel(const float* __restrict__ proto, const float* __restrict__ coef, int64_t npix, int m,
                                   float* __restrict__ out) {
  const int64_t total = npix * m;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(t % m);
    const int64_t pix = t / m;
    const float4* pp = reinterpret_cast<const float4*>(proto + pix * ISB_NUM_COEF);
    const float4* cc = reinterpret_cast<const float4*>(coef + (size_t)j * ISB_NUM_COEF);
    float acc = 0.0f;
#pragma unroll
    for (int q = 0; q < ISB_NUM_COEF / 4; ++q) {
      const float4 a = __ldg(pp + q), c = __ldg(cc + q);
      acc = fmaf(a.x, c.x, acc);
      acc = fmaf(a.y, c.y, acc);
      acc = fmaf(a.z, c.z, acc);
      acc = fmaf(a.w, c.w, acc);
    }
    out[t] = acc;
  }
}

static inline unsigned grid_for(int64_t n, int threads, int num_sms) {
  int64_t g = (n + threads - 1) / threads;
  int64_t cap = (int64_t)num_sms * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

int launch_decode_boxes(isb_handle* h, const float* pred, const float* anchors, int64_t n, float v0, float v1, float* out,
                        cudaStream_t s) {
  decode_boxes_kernel<<<grid_for(n, 256, h->num_sms), 256, 0, s>>>(
      reinterpret_cast<const float4*>(pred), reinterpret_cast<const float4*>(anchors), n, v0, v1,
      reinterpret_cast<float4*>(out));
  ISB_LAUNCH_CHECK(h, "decode_boxes_kernel");
  return ISB_OK;
}
int launch_jaccard(isb_handle* h, const float* a, const float* b, int batch, int na, int nb, int iscrowd, float* out,
                   cudaStream_t s) {
  jaccard_kernel<<<grid_for((int64_t)batch * na * nb, 256, h->num_sms), 256, 0, s>>>(
      reinterpret_cast<const float4*>(a), reinterpret_cast<const float4*>(b), batch, na, nb, iscrowd, out);
  ISB_LAUNCH_CHECK(h, "jaccard_kernel");
  return ISB_OK;
}
int launch_correct_boxes(isb_handle* h, const float* in, int64_t m, int out_h, int out_w, float* out, cudaStream_t s) {
  correct_boxes_kernel<<<grid_for(m, 128, h->num_sms), 128, 0, s>>>(reinterpret_cast<const float4*>(in), m,
                                                                    (float)out_w, (float)out_h,
                                                                    reinterpret_cast<float4*>(out));
  ISB_LAUNCH_CHECK(h, "correct_boxes_kernel");
  return ISB_OK;
}
int launch_crop(isb_handle* h, const float* masks, const float* boxes, int hgt, int wid, int m, float* out,
                cudaStream_t s) {
  crop_kernel<<<grid_for((int64_t)hgt * wid * m, 256, h->num_sms), 256, 0, s>>>(
      masks, reinterpret_cast<const float4*>(boxes), hgt, wid, m, out);
  ISB_LAUNCH_CHECK(h, "crop_kernel");
  return ISB_OK;
}
int launch_det_layout(isb_handle* h, const float* boxes, int m, int hgt, int wid, int proto_h, int proto_w, DetRec* out,
                      int32_t* total_out, cudaStream_t s) {
  det_layout_kernel<<<1, 1024, 0, s>>>(reinterpret_cast<const float4*>(boxes), m, hgt, wid, proto_h, proto_w, out, total_out);
  ISB_LAUNCH_CHECK(h, "det_layout_kernel");
  return ISB_OK;
}
int launch_row_argmax(isb_handle* h, const float* scores, int n, int ncols, int32_t* labels, cudaStream_t s) {
  row_argmax_kernel<<<grid_for(n, 128, h->num_sms), 128, 0, s>>>(scores, n, ncols, labels);
  ISB_LAUNCH_CHECK(h, "row_argmax_kernel");
  return ISB_OK;
}
int launch_trad_gather(isb_handle* h, const int64_t* keep, const int32_t* count, int max_det, const float* boxes,
                       const float* scores, const int32_t* labels, const float* coef, float* out_box, float* out_score,
                       int64_t* out_class, float* out_coef, int32_t* out_count, cudaStream_t s) {
  trad_gather_kernel<<<grid_for((int64_t)max_det * 8, 256, h->num_sms), 256, 0, s>>>(
      keep, count, max_det, reinterpret_cast<const float4*>(boxes), scores, labels, coef, reinterpret_cast<float4*>(out_box),
      out_score, out_class, out_coef, out_count);
  ISB_LAUNCH_CHECK(h, "trad_gather_kernel");
  return ISB_OK;
}
int launch_mask_logits(isb_handle* h, const float* proto, const float* coef, int64_t npix, int m, float* out,
                       cudaStream_t s) {
  mask_logits_kernel<<<grid_for(npix * m, 256, h->num_sms), 256, 0, s>>>(proto, coef, npix, m, out);
  ISB_LAUNCH_CHECK(h, "mask_logits_kernel");
  return ISB_OK;
}

}  // namespace isb
