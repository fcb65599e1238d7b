// Mask R-CNN element-wise / gather kernels: BoxCoder.decode_single, maskrcnn_inference,
// resize_boxes, paste_masks_in_image and (Multi-scale) RoIAlign.
#include "isb_common.cuh"

namespace isb {

// ------------------------------------------------------------------------------------------
// det_utils.py:224-267  decode_single: rel_codes [n, 4k], boxes [n, 4] -> [n, 4k]
// One thread per (row, k) box: float4 in, float4 out, coalesced.  HBM-bound (48 B per box).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 rcnn_decode(float4 rel, float4 b, float wx, float wy, float ww, float wh, float clip) {
  const float width = b.z - b.x, height = b.w - b.y;
  const float ctr_x = b.x + 0.5f * width, ctr_y = b.y + 0.5f * height;
  const float dx = rel.x / wx, dy = rel.y / wy;
  const float dw = fminf(rel.z / ww, clip), dh = fminf(rel.w / wh, clip);
  const float pcx = dx * width + ctr_x, pcy = dy * height + ctr_y;
  const float pw = expf(dw) * width, ph = expf(dh) * height;
  return make_float4(pcx - 0.5f * pw, pcy - 0.5f * ph, pcx + 0.5f * pw, pcy + 0.5f * ph);
}

__global__ void box_decode_kernel(const float4* __restrict__ rel, const float4* __restrict__ boxes, int64_t n, int k,
                                  float wx, float wy, float ww, float wh, float clip, float4* __restrict__ out) {
  const int64_t total = n * k;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x)
    out[t] = rcnn_decode(rel[t], __ldg(boxes + t / k), wx, wy, ww, wh, clip);
}

// ------------------------------------------------------------------------------------------
// roi_head.py:59-91  maskrcnn_inference: sigmoid of the channel picked by the label only
// (the reference computes all C channels, then selects).
// ------------------------------------------------------------------------------------------
__global__ void mask_select_sigmoid_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, int m,
                                           int num_classes, int plane, float* __restrict__ out) {
  const int64_t total = (int64_t)m * plane;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(t / plane), s = (int)(t - (int64_t)i * plane);
    const int64_t c = labels[i];
    out[t] = sigmoidf_ref(logits[((int64_t)i * num_classes + c) * plane + s]);
  }
}

// transform.py:462-484
__global__ void resize_boxes_kernel(const float4* __restrict__ in, int64_t m, float rw, float rh, float4* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 b = in[i];
    out[i] = make_float4(b.x * rw, b.y * rh, b.z * rw, b.w * rh);
  }
}

// ------------------------------------------------------------------------------------------
// transform.py:212-230  paste_masks_in_image (+ expand_masks :171, expand_boxes :145,
// paste_mask_in_image :182).  Output float32 [m,1,H,W] written exactly once with 16-byte stores:
// HBM-write-bound (427 MB at m=100, 800x1333).  A thread produces 4 consecutive output floats; each
// one is either 0 (outside the integer box) or a bilinear sample of the zero-padded SxS mask
// resized to the box (aten align_corners=False arithmetic), read through L1.
// ------------------------------------------------------------------------------------------
struct PasteBox { int x0, y0, x1, y1; };

__device__ __forceinline__ PasteBox paste_box(float4 b, float scale) {
  // expand_boxes, then .to(int64) (truncation toward zero)
  float w_half = (b.z - b.x) * 0.5f, h_half = (b.w - b.y) * 0.5f;
  const float x_c = (b.z + b.x) * 0.5f, y_c = (b.w + b.y) * 0.5f;
  w_half *= scale;
  h_half *= scale;
  PasteBox r;
  r.x0 = (int)(x_c - w_half);
  r.x1 = (int)(x_c + w_half);
  r.y0 = (int)(y_c - h_half);
  r.y1 = (int)(y_c + h_half);
  return r;
}

__device__ __forceinline__ float padded_at(const float* __restrict__ mk, int S, int pad, int y, int x) {
  const int yy = y - pad, xx = x - pad;
  return (yy >= 0 && yy < S && xx >= 0 && xx < S) ? __ldg(mk + yy * S + xx) : 0.0f;
}

__global__ void __launch_bounds__(256) paste_masks_kernel(const float* __restrict__ masks, const float4* __restrict__ boxes,
                                                          int m, int S, int pad, int im_h, int im_w,
                                                          float* __restrict__ out) {
  const int64_t total = (int64_t)m * im_h * im_w;
  const int64_t nvec = (total + 3) >> 2;
  const int Sp = S + 2 * pad;
  const float scale = (float)((double)Sp / (double)S);  // transform.py:178, cast to the tensor dtype at :156
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e0 = v << 2;
    int64_t pix = e0 % ((int64_t)im_h * im_w);
    int i = (int)(e0 / ((int64_t)im_h * im_w));
    int y = (int)(pix / im_w), x = (int)(pix - (int64_t)y * im_w);
    float vals[4];
    int cur = -1;
    PasteBox pb = {0, 0, 0, 0};
    int w = 1, hgt = 1;
    float sx = 0.f, sy = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float r = 0.0f;
      if (e0 + k < total) {
        if (i != cur) {
          pb = paste_box(__ldg(boxes + i), scale);
          w = max(pb.x1 - pb.x0 + 1, 1);
          hgt = max(pb.y1 - pb.y0 + 1, 1);
          sx = (float)Sp / (float)w;
          sy = (float)Sp / (float)hgt;
          cur = i;
        }
        // transform.py:202-208: the pasted window is [max(x0,0), min(x1+1, W)) x [max(y0,0), min(y1+1, H))
        if (x >= pb.x0 && x <= pb.x1 && y >= pb.y0 && y <= pb.y1) {
          int iy0, iy1, ix0, ix1;
          float ly0, ly1, lx0, lx1;
          bilinear_src(y - pb.y0, sy, Sp, iy0, iy1, ly0, ly1);
          bilinear_src(x - pb.x0, sx, Sp, ix0, ix1, lx0, lx1);
          const float* mk = masks + (int64_t)i * S * S;
          const float top = lx0 * padded_at(mk, S, pad, iy0, ix0) + lx1 * padded_at(mk, S, pad, iy0, ix1);
          const float bot = lx0 * padded_at(mk, S, pad, iy1, ix0) + lx1 * padded_at(mk, S, pad, iy1, ix1);
          r = ly0 * top + ly1 * bot;
        }
        if (++x == im_w) { x = 0; if (++y == im_h) { y = 0; ++i; } }
      }
      vals[k] = r;
    }
    if (e0 + 3 < total) {
      reinterpret_cast<float4*>(out)[v] = make_float4(vals[0], vals[1], vals[2], vals[3]);
    } else {
      for (int k = 0; k < 4 && e0 + k < total; ++k) out[e0 + k] = vals[k];
    }
  }
}

// ------------------------------------------------------------------------------------------
// torchvision MultiScaleRoIAlign (ctor sites faster_rcnn_framework.py:326-329, mask_rcnn.py:150;
// call sites roi_head.py:491,540): FPN level = floor(k0 + log2(sqrt(area)/s0) + eps) clamped to
// [k_min,k_max]; per-level roi_align(aligned=False, sampling_ratio>0); result in RoI order.
// One CTA per RoI: the (y, x) sample tables (index pair, weights, validity) are computed once in
// shared memory and reused by all C channels; outputs are written coalesced [K,C,oh,ow].
// ------------------------------------------------------------------------------------------
struct AxisSample { int lo, hi; float l, h; int bad; };

__device__ __forceinline__ AxisSample axis_sample(float v, int size) {
  AxisSample s;
  s.bad = (v < -1.0f || v > (float)size) ? 1 : 0;
  if (v <= 0.0f) v = 0.0f;
  int lo = (int)v, hi;
  if (lo >= size - 1) { hi = lo = size - 1; v = (float)lo; } else { hi = lo + 1; }
  s.lo = lo; s.hi = hi;
  s.l = v - (float)lo;
  s.h = 1.0f - s.l;
  return s;
}

constexpr int kRoiThreads = 256;
constexpr int kRoiAxisMax = 128;  // out * sampling_ratio per axis

__global__ void __launch_bounds__(kRoiThreads) ms_roi_align_kernel(RoiLevels L, const float4* __restrict__ boxes,
                                                                   const int32_t* __restrict__ batch_idx, int channels,
                                                                   int oh, int ow, int sr, float* __restrict__ out) {
  __shared__ AxisSample ys[kRoiAxisMax], xs[kRoiAxisMax];
  __shared__ int s_level;
  const int r = blockIdx.x, tid = threadIdx.x;
  const float4 b = boxes[r];
  if (tid == 0) {
    int lvl = 0;
    if (L.num_levels > 1) {
      const float s = sqrtf((b.z - b.x) * (b.w - b.y));
      float t = floorf(L.canonical_level + log2f(s / L.canonical_scale) + L.eps);
      t = fminf(fmaxf(t, (float)L.k_min), (float)L.k_max);
      lvl = (int)(long long)t - L.k_min;
    }
    s_level = lvl;
  }
  __syncthreads();
  const int lvl = s_level;
  const int fh = L.h[lvl], fw = L.w[lvl];
  const float ss = L.scale[lvl];
  const float start_w = b.x * ss, start_h = b.y * ss, end_w = b.z * ss, end_h = b.w * ss;
  const float roi_w = fmaxf(end_w - start_w, 1.0f), roi_h = fmaxf(end_h - start_h, 1.0f);
  const float bin_h = roi_h / (float)oh, bin_w = roi_w / (float)ow;
  for (int t = tid; t < oh * sr; t += kRoiThreads) {
    const int ph = t / sr, iy = t - ph * sr;
    ys[t] = axis_sample(start_h + (float)ph * bin_h + (((float)iy + 0.5f) * bin_h) / (float)sr, fh);
  }
  for (int t = tid; t < ow * sr; t += kRoiThreads) {
    const int pw = t / sr, ix = t - pw * sr;
    xs[t] = axis_sample(start_w + (float)pw * bin_w + (((float)ix + 0.5f) * bin_w) / (float)sr, fw);
  }
  __syncthreads();
  const float count = (float)max(sr * sr, 1);
  const int bins = oh * ow;
  const int img = batch_idx ? batch_idx[r] : 0;
  const float* fbase = L.feat[lvl] + (size_t)img * channels * fh * fw;
  float* obase = out + (size_t)r * channels * bins;
  for (int o = tid; o < channels * bins; o += kRoiThreads) {
    const int c = o / bins, bin = o - c * bins;
    const int ph = bin / ow, pw = bin - ph * ow;
    const float* plane = fbase + (size_t)c * fh * fw;
    float acc = 0.0f;
    for (int iy = 0; iy < sr; ++iy) {
      const AxisSample sy = ys[ph * sr + iy];
      for (int ix = 0; ix < sr; ++ix) {
        const AxisSample sx = xs[pw * sr + ix];
        float val = 0.0f;
        if (!(sy.bad | sx.bad)) {
          const float w1 = sy.h * sx.h, w2 = sy.h * sx.l, w3 = sy.l * sx.h, w4 = sy.l * sx.l;
          const float v1 = __ldg(plane + sy.lo * fw + sx.lo), v2 = __ldg(plane + sy.lo * fw + sx.hi);
          const float v3 = __ldg(plane + sy.hi * fw + sx.lo), v4 = __ldg(plane + sy.hi * fw + sx.hi);
          val = ((w1 * v1 + w2 * v2) + w3 * v3) + w4 * v4;
        }
        acc += val;
      }
    }
    obase[o] = acc / count;
  }
}

// ---------------------------------------------------------------------------- launchers
static inline unsigned grid_for(int64_t n, int threads, int num_sms, int per_sm = 16) {
  int64_t g = (n + threads - 1) / threads;
  const int64_t cap = (int64_t)num_sms * per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

int launch_box_decode(isb_handle* h, const float* rel, const float* boxes, int64_t n, int k, const float w[4], float clip,
                      float* out, cudaStream_t s) {
  box_decode_kernel<<<grid_for(n * k, 256, h->num_sms), 256, 0, s>>>(
      reinterpret_cast<const float4*>(rel), reinterpret_cast<const float4*>(boxes), n, k, w[0], w[1], w[2], w[3], clip,
      reinterpret_cast<float4*>(out));
  ISB_LAUNCH_CHECK(h, "box_decode_kernel");
  return ISB_OK;
}
int launch_mask_select_sigmoid(isb_handle* h, const float* logits, const int64_t* labels, int m, int num_classes, int plane,
                               float* out, cudaStream_t s) {
  mask_select_sigmoid_kernel<<<grid_for((int64_t)m * plane, 256, h->num_sms), 256, 0, s>>>(logits, labels, m, num_classes,
                                                                                           plane, out);
  ISB_LAUNCH_CHECK(h, "mask_select_sigmoid_kernel");
  return ISB_OK;
}
int launch_resize_boxes(isb_handle* h, const float* in, int64_t m, float rw, float rh, float* out, cudaStream_t s) {
  resize_boxes_kernel<<<grid_for(m, 128, h->num_sms), 128, 0, s>>>(reinterpret_cast<const float4*>(in), m, rw, rh,
                                                                   reinterpret_cast<float4*>(out));
  ISB_LAUNCH_CHECK(h, "resize_boxes_kernel");
  return ISB_OK;
}
int launch_paste_masks(isb_handle* h, const float* masks, const float* boxes, int m, int S, int pad, int im_h, int im_w,
                       float* out, cudaStream_t s) {
  const int64_t nvec = ((int64_t)m * im_h * im_w + 3) / 4;
  paste_masks_kernel<<<grid_for(nvec, 256, h->num_sms, 32), 256, 0, s>>>(masks, reinterpret_cast<const float4*>(boxes), m, S,
                                                                         pad, im_h, im_w, out);
  ISB_LAUNCH_CHECK(h, "paste_masks_kernel");
  return ISB_OK;
}
int launch_ms_roi_align(isb_handle* h, const RoiLevels& L, const float* boxes, const int32_t* batch_idx, int num_rois,
                        int channels, int oh, int ow, int sr, float* out, cudaStream_t s) {
  ms_roi_align_kernel<<<num_rois, kRoiThreads, 0, s>>>(L, reinterpret_cast<const float4*>(boxes), batch_idx, channels, oh,
                                                       ow, sr, out);
  ISB_LAUNCH_CHECK(h, "ms_roi_align_kernel");
  return ISB_OK;
}

}  // namespace isb
