// Mask R-CNN element-wise / gather kernels: BoxCoder.decode_single, maskrcnn_inference,
// resize_boxes, paste_masks_in_image and (Multi-scale) RoIAlign.
#include <stdlib.h>

#include <cuda.h>
#include <mutex>
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
// paste_mask_in_image :182).  Output float32 [m,1,H,W] written exactly once: HBM-write-bound (427 MB at
// m=100, 800x1333).  Each element is either 0 (outside the integer box) or a bilinear sample of the
// zero-padded SxS mask resized to the box (aten align_corners=False arithmetic), read through L1.
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

// One pasted value: bilinear sample of the zero-padded mask resized to the integer box (transform.py:182-209).
struct PasteWin {
  PasteBox pb;
  float sx, sy;
  const float* mk;
};
__device__ __forceinline__ PasteWin paste_win(const float* __restrict__ masks, const float4* __restrict__ boxes, int i, int S,
                                              int Sp, float scale) {
  PasteWin w;
  w.pb = paste_box(__ldg(boxes + i), scale);
  w.sx = (float)Sp / (float)max(w.pb.x1 - w.pb.x0 + 1, 1);
  w.sy = (float)Sp / (float)max(w.pb.y1 - w.pb.y0 + 1, 1);
  w.mk = masks + (int64_t)i * S * S;
  return w;
}
__device__ __forceinline__ float paste_sample(const PasteWin& w, int S, int pad, int Sp, int y, int x) {
  int iy0, iy1, ix0, ix1;
  float ly0, ly1, lx0, lx1;
  bilinear_src(y - w.pb.y0, w.sy, Sp, iy0, iy1, ly0, ly1);
  bilinear_src(x - w.pb.x0, w.sx, Sp, ix0, ix1, lx0, lx1);
  const float top = lx0 * padded_at(w.mk, S, pad, iy0, ix0) + lx1 * padded_at(w.mk, S, pad, iy0, ix1);
  const float bot = lx0 * padded_at(w.mk, S, pad, iy1, ix0) + lx1 * padded_at(w.mk, S, pad, iy1, ix1);
  return ly0 * top + ly1 * bot;
}

// The output is one flat stream of m*H*W floats cut into 4 KB chunks; chunk c belongs to warp (c mod #warps), so
// neighbouring chunks (the rows of one box) spread over the whole grid.  A chunk whose rows miss its mask's box -- most
// of them -- leaves as ONE bulk asynchronous store from a shared all-zero buffer (cp.async.bulk.global.shared::cta, SASS
// UBLKCP): one lane issues it and no SM instruction touches the data.  Any other chunk is composed in one of the warp's
// two private shared-memory buffers -- cleared with 16-byte stores, then only the columns inside the box are sampled,
// row piece by row piece: the row's bilinear taps are warp-uniform, the two padded source rows sit in registers (one
// column per lane) and the four taps of a pixel are warp shuffles -- and leaves the same way; the warp only waits
// (wait_group.read) before it overwrites a buffer.  No CTA barrier after the prologue.
constexpr int kPasteChunk = 1024;  // floats per bulk store
constexpr int kPasteWarps = 16;
constexpr size_t kPasteSmem = (size_t)(1 + 2 * kPasteWarps) * kPasteChunk * sizeof(float);

template <int N>
__device__ __forceinline__ void bulk_wait_read_n() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

__global__ void __launch_bounds__(kPasteWarps * 32, 1)
paste_masks_bulk_kernel(const float* __restrict__ masks, const float4* __restrict__ boxes, int m, int S, int pad, int im_h,
                        int im_w, float* __restrict__ out) {
  extern __shared__ __align__(128) float s_paste[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* s_zero = s_paste;
  float* s_buf = s_paste + kPasteChunk + (size_t)warp * 2 * kPasteChunk;
  for (int t = threadIdx.x; t < kPasteChunk; t += blockDim.x) s_zero[t] = 0.0f;
  fence_proxy_async();
  __syncthreads();
  const unsigned long long plane = (unsigned long long)im_h * im_w, total = (unsigned long long)m * plane;
  const unsigned long long nchunks = (total + kPasteChunk - 1) / kPasteChunk;
  const int Sp = S + 2 * pad;
  const float scale = (float)((double)Sp / (double)S);  // transform.py:178, cast to the tensor dtype at :156
  int flip = 0;
  for (unsigned long long c = (unsigned long long)blockIdx.x * kPasteWarps + warp; c < nchunks;
       c += (unsigned long long)gridDim.x * kPasteWarps) {
    const unsigned long long e0 = c * kPasteChunk;
    const int n = (int)min((unsigned long long)kPasteChunk, total - e0);
    const int nvec = n & ~3;
    // walk the row pieces of the chunk (warp-uniform): (mask ii, row y, columns [x, x + len))
    int ii = (int)(e0 / plane);
    const unsigned long long pix0 = e0 - (unsigned long long)ii * plane;
    int y = (int)(pix0 / (unsigned)im_w), x = (int)(pix0 - (unsigned long long)y * im_w);
    PasteWin w = paste_win(masks, boxes, ii, S, Sp, scale);
    float* buf = nullptr;
    for (int off = 0; off < n;) {
      const int len = min(im_w - x, n - off);
      // transform.py:202-208: the pasted window is [max(x0,0), min(x1+1, W)) x [max(y0,0), min(y1+1, H))
      const int lo = max(x, w.pb.x0), hi = min(x + len - 1, w.pb.x1);
      if (y >= w.pb.y0 && y <= w.pb.y1 && lo <= hi) {
        if (!buf) {  // first piece of the chunk that meets its box: take a buffer and clear it
          buf = s_buf + flip * kPasteChunk;
          flip ^= 1;
          if (lane == 0) bulk_wait_read1();
          __syncwarp();
#pragma unroll
          for (int k = 0; k < kPasteChunk / 128; ++k)
            reinterpret_cast<float4*>(buf)[lane + 32 * k] = make_float4(0.f, 0.f, 0.f, 0.f);
          __syncwarp();
        }
        int iy0, iy1;
        float ly0, ly1;
        bilinear_src(y - w.pb.y0, w.sy, Sp, iy0, iy1, ly0, ly1);
        float* dst = buf + off - x;  // dst[column]
        if (Sp <= 32) {
          // lane l keeps column l of the two padded source rows; the four taps of a pixel are shuffles
          const float p0 = lane < Sp ? padded_at(w.mk, S, pad, iy0, lane) : 0.0f;
          const float p1 = lane < Sp ? padded_at(w.mk, S, pad, iy1, lane) : 0.0f;
          for (int xb = lo; xb <= hi; xb += 32) {
            const int xx = xb + lane;
            int ix0, ix1;
            float lx0, lx1;
            bilinear_src(xx - w.pb.x0, w.sx, Sp, ix0, ix1, lx0, lx1);
            const float top = lx0 * __shfl_sync(0xffffffffu, p0, ix0) + lx1 * __shfl_sync(0xffffffffu, p0, ix1);
            const float bot = lx0 * __shfl_sync(0xffffffffu, p1, ix0) + lx1 * __shfl_sync(0xffffffffu, p1, ix1);
            if (xx <= hi) dst[xx] = ly0 * top + ly1 * bot;
          }
        } else {
          for (int xx = lo + lane; xx <= hi; xx += 32) dst[xx] = paste_sample(w, S, pad, Sp, y, xx);
        }
      }
      off += len;
      x = 0;
      if (++y == im_h) {
        y = 0;
        if (++ii < m) w = paste_win(masks, boxes, ii, S, Sp, scale);
      }
    }
    const float* src = s_zero;
    if (buf) {
      fence_proxy_async();
      __syncwarp();
      src = buf;
    } else if (lane == 0) {
      bulk_wait_read_n<6>();
    }
    if (lane == 0 && nvec > 0) {
      bulk_s2g(out + e0, src, (uint32_t)nvec * sizeof(float));
      bulk_commit();
    }
    if (lane < n - nvec) out[e0 + nvec + lane] = src[nvec + lane];  // < 4 trailing floats of the whole output
  }
  if (lane == 0) bulk_wait0();
}

// ------------------------------------------------------------------------------------------
// Wrapper tail inst_model/Mask_RCNN/mask_rcnn.py:174-198 fused with the paste (transform.py:212-230): score filter,
// paste into (paste_h, paste_w), bilinear resize to the original image size, arg-max over detections, > 0.5, class
// image.  None of the intermediates ([m,1,Hp,Wp] floats: 118 MB at 544^2 x 100; [H,W,m] at the original size) exists:
// a thread owns one output pixel, finds its four source taps in the pasted frame, and for each kept detection whose
// pasted window contains a tap evaluates the paste at those taps on the fly (same arithmetic as paste_masks_kernel).
// ------------------------------------------------------------------------------------------
struct PasteDet {
  int x0, y0, x1, y1;  // integer box after expand + truncation
  float sx, sy;        // padded mask size / box size
  int src;             // row of masks / boxes
  int cls;             // label
};

__device__ __forceinline__ float pasted_at(const float* __restrict__ mk, const PasteDet& d, int S, int pad, int Sp, int y, int x) {
  if (x < d.x0 || x > d.x1 || y < d.y0 || y > d.y1) return 0.0f;
  int iy0, iy1, ix0, ix1;
  float ly0, ly1, lx0, lx1;
  bilinear_src(y - d.y0, d.sy, Sp, iy0, iy1, ly0, ly1);
  bilinear_src(x - d.x0, d.sx, Sp, ix0, ix1, lx0, lx1);
  const float top = lx0 * padded_at(mk, S, pad, iy0, ix0) + lx1 * padded_at(mk, S, pad, iy0, ix1);
  const float bot = lx0 * padded_at(mk, S, pad, iy1, ix0) + lx1 * padded_at(mk, S, pad, iy1, ix1);
  return ly0 * top + ly1 * bot;
}

constexpr int kPcmThreads = 256;

__global__ void __launch_bounds__(kPcmThreads)
paste_class_map_kernel(const float* __restrict__ masks, const float4* __restrict__ boxes, const int64_t* __restrict__ labels,
                       const float* __restrict__ scores, int m, int S, int pad, int paste_h, int paste_w, int out_h, int out_w,
                       float confidence, uint8_t* __restrict__ class_map, int64_t* __restrict__ masks_arg,
                       int32_t* __restrict__ kept_count) {
  extern __shared__ __align__(16) unsigned char pcm_smem[];
  PasteDet* dets = reinterpret_cast<PasteDet*>(pcm_smem);
  __shared__ int s_wcount[kPcmThreads / 32];
  __shared__ int s_n;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int Sp = S + 2 * pad;
  const float scale = (float)((double)Sp / (double)S);
  // kept detections (score > confidence, mask_rcnn.py:174), in their original order
  if (tid == 0) s_n = 0;
  __syncthreads();
  for (int base = 0; base < m; base += kPcmThreads) {
    const int i = base + tid;
    const bool keep = i < m && scores[i] > confidence;
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_wcount[warp] = __popc(bal);
    __syncthreads();
    int off = s_n, tot = 0;
    for (int w = 0; w < kPcmThreads / 32; ++w) {
      if (w < warp) off += s_wcount[w];
      tot += s_wcount[w];
    }
    if (keep) {
      const PasteBox pb = paste_box(boxes[i], scale);
      PasteDet d;
      d.x0 = pb.x0; d.y0 = pb.y0; d.x1 = pb.x1; d.y1 = pb.y1;
      d.sx = (float)Sp / (float)max(pb.x1 - pb.x0 + 1, 1);
      d.sy = (float)Sp / (float)max(pb.y1 - pb.y0 + 1, 1);
      d.src = i;
      d.cls = (int)labels[i];
      dets[off + __popc(bal & ((1u << lane) - 1u))] = d;
    }
    __syncthreads();
    if (tid == 0) s_n += tot;
    __syncthreads();
  }
  const int n = s_n;
  if (blockIdx.x == 0 && tid == 0 && kept_count) *kept_count = n;
  const float ry = (float)paste_h / (float)out_h, rx = (float)paste_w / (float)out_w;  // F.interpolate, align_corners=False
  const int total = out_h * out_w;
  for (int q = blockIdx.x * kPcmThreads + tid; q < total; q += gridDim.x * kPcmThreads) {
    const int y = q / out_w, x = q - y * out_w;
    int iy0, iy1, ix0, ix1;
    float ly0, ly1, lx0, lx1;
    bilinear_src(y, ry, paste_h, iy0, iy1, ly0, ly1);
    bilinear_src(x, rx, paste_w, ix0, ix1, lx0, lx1);
    float best = 0.0f;
    int arg = 0;
    for (int k = 0; k < n; ++k) {
      const PasteDet d = dets[k];
      if (iy1 < d.y0 || iy0 > d.y1 || ix1 < d.x0 || ix0 > d.x1) continue;  // all four taps are zeros of the paste
      const float* mk = masks + (int64_t)d.src * S * S;
      const float top = lx0 * pasted_at(mk, d, S, pad, Sp, iy0, ix0) + lx1 * pasted_at(mk, d, S, pad, Sp, iy0, ix1);
      const float bot = lx0 * pasted_at(mk, d, S, pad, Sp, iy1, ix0) + lx1 * pasted_at(mk, d, S, pad, Sp, iy1, ix1);
      const float v = ly0 * top + ly1 * bot;
      if (v > best) { best = v; arg = k; }
    }
    class_map[q] = (n > 0 && best > 0.5f) ? (uint8_t)(dets[arg].cls + 1) : (uint8_t)0;
    if (masks_arg) masks_arg[q] = arg;
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
constexpr int kRoiAxisMax = 128;      // out * sampling_ratio per axis
constexpr int kRoiTileFloats = 8 * 1024;   // tap tiles of one pass of channels + the pass's output staging (32 KB; swept 6-18 K)
constexpr int kRoiMaxEntries = 1024;       // (bin, sample) table entries the launcher will size shared memory for (32 KB)

__device__ __forceinline__ void cp_async4(uint32_t dst_smem, const void* src_gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst_smem), "l"(src_gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// sample of one axis in COMPACT coordinates: positions of its two taps in the list of distinct rows (columns)
// the RoI touches, interpolation weight, validity
struct TapSample { int32_t idx; float l; };  // idx = lo | hi << 8 | bad << 16

// One thread: distinct tap coordinates of an axis, in ascending order, and every sample's positions in that list.
// The samples are monotone along the axis, so a new coordinate is either one of the last two listed or larger.
__device__ void build_tap_list(const AxisSample* __restrict__ smp, int n, int* __restrict__ list, int* __restrict__ n_list,
                               TapSample* __restrict__ out) {
  int cnt = 0;
  for (int t = 0; t < n; ++t) {
    const AxisSample a = smp[t];
    TapSample ts;
    ts.l = a.l;
    ts.idx = 1 << 16;
    if (!a.bad) {
      int pos[2];
      const int v[2] = {a.lo, a.hi};
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        if (cnt >= 1 && list[cnt - 1] == v[k]) pos[k] = cnt - 1;
        else if (cnt >= 2 && list[cnt - 2] == v[k]) pos[k] = cnt - 2;
        else { list[cnt] = v[k]; pos[k] = cnt++; }
      }
      ts.idx = pos[0] | (pos[1] << 8);
    }
    out[t] = ts;
  }
  *n_list = cnt;
}

// grid (RoI, channel group).  The CTA works out the RoI's FPN level and its (y, x) sample tables once, then the
// lists of DISTINCT feature rows and columns those samples touch (at most 2 * out * sampling_ratio each, whatever
// the RoI's size or aspect ratio) and, per (bin, sample), the four tap positions inside a channel tile and their
// weights.  The group's channels go through shared memory in passes of up to 32: the ny x nx tap tile of every
// channel of the pass is gathered with asynchronous copies (cp.async, coalesced along the feature rows; a thread owns
// a tap and copies it for all channels of the pass), stored channel by channel with an ODD pitch.  In the reduction
// a lane is a CHANNEL and a warp walks output bins: tap positions and weights are warp-uniform (two broadcast
// 16-byte reads per sample), the 16 taps of a bin are 16 conflict-free shared-memory reads (lane stride = the odd
// pitch), and there is no per-lane index arithmetic.  The first version,
// lane = output element, spent 136 M warp instructions and 36 M shared-memory wavefronts (13 M of them bank
// conflicts) on 1000 RoIs (ncu); this one needs about a quarter of both.  Results are staged [channel][bin] in shared
// memory and leave as one contiguous, coalesced run per pass (the channels of a RoI are adjacent in [K,C,oh,ow]),
// while the next pass's gather is already in flight.  TMA tensor tiles are not usable in general here: the row
// pitch of an FPN level need not be a multiple of 16 bytes (42 floats at the coarsest level of an 800 x 1344 input).
// SR > 0: sampling ratio known at compile time (the reference always passes 2); SR == 0: run-time value.
template <int SR>
__global__ void __launch_bounds__(kRoiThreads, 4) ms_roi_align_kernel(RoiLevels L, const float4* __restrict__ boxes,
                                                                   const int32_t* __restrict__ batch_idx, int channels,
                                                                   int oh, int ow, int sr_rt, int cpc, int tile_floats, int table_cap,
                                                                      float* __restrict__ out) {
  const int sr = SR > 0 ? SR : sr_rt;
  extern __shared__ __align__(16) float s_buf[];  // [kRoiTileFloats] tiles + output staging | (bin, sample) table
  __shared__ AxisSample ys[kRoiAxisMax], xs[kRoiAxisMax];
  __shared__ TapSample tys[kRoiAxisMax], txs[kRoiAxisMax];
  __shared__ int ylist[2 * kRoiAxisMax], xlist[2 * kRoiAxisMax];
  __shared__ int s_level, s_ny, s_nx;
  int4* s_off = reinterpret_cast<int4*>(s_buf + tile_floats);          // tap offsets of a sample inside a channel tile
  float4* s_wt = reinterpret_cast<float4*>(s_off + table_cap);         // its four bilinear weights
  const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cbase = blockIdx.y * cpc, cend = min(channels, cbase + cpc);
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
  const int ny = oh * sr, nx = ow * sr;
  for (int t = tid; t < ny; t += kRoiThreads) {
    const int ph = t / sr, iy = t - ph * sr;
    ys[t] = axis_sample(start_h + (float)ph * bin_h + (((float)iy + 0.5f) * bin_h) / (float)sr, fh);
  }
  for (int t = tid; t < nx; t += kRoiThreads) {
    const int pw = t / sr, ix = t - pw * sr;
    xs[t] = axis_sample(start_w + (float)pw * bin_w + (((float)ix + 0.5f) * bin_w) / (float)sr, fw);
  }
  __syncthreads();
  if (tid == 0) build_tap_list(ys, ny, ylist, &s_ny, tys);
  if (tid == 32) build_tap_list(xs, nx, xlist, &s_nx, txs);
  __syncthreads();
  const int nyl = s_ny, nxl = s_nx, plane = nyl * nxl;
  const int nsmp = sr * sr, entries = oh * ow * nsmp;
  const float count = SR > 0 ? (float)(SR * SR) : (float)max(nsmp, 1);
  const int bins = oh * ow;
  const int img = batch_idx ? batch_idx[r] : 0;
  const float* fbase = L.feat[lvl] + (size_t)img * channels * fh * fw;
  float* obase = out + (size_t)r * channels * bins;
  if (plane == 0) {  // every sample falls outside the map: zeros
    for (int o = cbase * bins + tid; o < cend * bins; o += kRoiThreads) obase[o] = 0.0f;
    return;
  }
  const int pitch = plane | 1;
  if (entries > table_cap || nyl > 255 || nxl > 255 || pitch + bins > tile_floats) {  // very large out * sampling_ratio: direct gather
    for (int o = cbase * bins + tid; o < cend * bins; o += kRoiThreads) {
      const int c = o / bins, bin = o - c * bins;
      const int ph = bin / ow, pw = bin - ph * ow;
      const float* pl = fbase + (size_t)c * fh * fw;
      float acc = 0.0f;
      for (int iy = 0; iy < sr; ++iy) {
        const AxisSample sy = ys[ph * sr + iy];
        for (int ix = 0; ix < sr; ++ix) {
          const AxisSample sx = xs[pw * sr + ix];
          float val = 0.0f;
          if (!(sy.bad | sx.bad)) {
            const float w1 = sy.h * sx.h, w2 = sy.h * sx.l, w3 = sy.l * sx.h, w4 = sy.l * sx.l;
            const float v1 = __ldg(pl + sy.lo * fw + sx.lo), v2 = __ldg(pl + sy.lo * fw + sx.hi);
            const float v3 = __ldg(pl + sy.hi * fw + sx.lo), v4 = __ldg(pl + sy.hi * fw + sx.hi);
            val = ((w1 * v1 + w2 * v2) + w3 * v3) + w4 * v4;
          }
          acc += val;
        }
      }
      obase[o] = acc / count;
    }
    return;
  }
  // (bin, sample) table: where the sample's four taps sit inside a channel tile and their weights w1..w4
  // (roi_align's hy*hx, hy*lx, ly*hx, ly*lx).  A sample outside the map contributes 0 in the reference; here it
  // gets four zero weights on tap 0, which is the same for finite features.
  for (int e = tid; e < entries; e += kRoiThreads) {
    const int bin = e / nsmp, k = e - bin * nsmp, iy = k / sr, ix = k - iy * sr;
    const int ph = bin / ow, pw = bin - ph * ow;
    const TapSample sy = tys[ph * sr + iy], sx = txs[pw * sr + ix];
    int4 o = make_int4(0, 0, 0, 0);
    float4 wq = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!((sy.idx | sx.idx) >> 16)) {
      const float syl = sy.l, syh = 1.0f - syl, sxl = sx.l, sxh = 1.0f - sxl;
      const int rlo = (sy.idx & 0xff) * nxl, rhi = ((sy.idx >> 8) & 0xff) * nxl;
      const int xlo = sx.idx & 0xff, xhi = (sx.idx >> 8) & 0xff;
      o = make_int4(rlo + xlo, rlo + xhi, rhi + xlo, rhi + xhi);
      wq = make_float4(syh * sxh, syh * sxl, syl * sxh, syl * sxl);
    }
    s_off[e] = o;
    s_wt[e] = wq;
  }
  // channels per pass: the largest power of two <= 32 whose tiles and outputs fit
  int cg = 32;
  while (cg > 1 && cg * (pitch + bins) > tile_floats) cg >>= 1;
  float* otile = s_buf + cg * pitch;
  const uint32_t sbase = smem_u32(s_buf);
  const int ch = lane & (cg - 1);                                  // lane -> (channel of the pass, bin slot)
  const int slots = kRoiThreads / cg, slot = tid / cg;             // bin slots of the CTA (8 when cg == 32)
  const float* tc = s_buf + ch * pitch;
  const size_t fhw = (size_t)fh * fw;
  // a thread owns taps tid, tid + 256, ... of the tile and copies them for every channel of the pass: the tap's global
  // offset is worked out once, then it is one cp.async and two pointer bumps per element
#define ROI_STAGE(C0)                                                                    \
  do {                                                                                   \
    const int nc_ = min(cg, cend - (C0));                                                \
    const float* src0_ = fbase + (size_t)(C0) * fhw;                                     \
    for (int i = tid; i < plane; i += kRoiThreads) {                                     \
      const int yi = i / nxl, xi = i - yi * nxl;                                         \
      const float* src_ = src0_ + (ylist[yi] * fw + xlist[xi]);                          \
      uint32_t d_ = sbase + 4u * i;                                                      \
      _Pragma("unroll 4") for (int c = 0; c < nc_; ++c) {                                \
        cp_async4(d_, src_);                                                             \
        src_ += fhw;                                                                     \
        d_ += 4u * pitch;                                                                \
      }                                                                                  \
    }                                                                                    \
    cp_async_commit();                                                                   \
  } while (0)
  ROI_STAGE(cbase);
  for (int c0 = cbase; c0 < cend; c0 += cg) {
    const int nc = min(cg, cend - c0);
    cp_async_wait0();
    __syncthreads();  // tiles of this pass (and, the first time, the sample table) complete; otile is free again
    if (ch < nc) {
      for (int bin = slot; bin < bins; bin += slots) {
        float acc = 0.0f;
        const int e0 = bin * nsmp;
#pragma unroll
        for (int k = 0; k < (SR > 0 ? SR * SR : nsmp); ++k) {
          const int4 o = s_off[e0 + k];
          const float4 wq = s_wt[e0 + k];
          acc += ((wq.x * tc[o.x] + wq.y * tc[o.y]) + wq.z * tc[o.z]) + wq.w * tc[o.w];
        }
        otile[ch * bins + bin] = acc / count;
      }
    }
    __syncthreads();  // tiles consumed, outputs staged
    if (c0 + cg < cend) ROI_STAGE(c0 + cg);
    float* dst = obase + (size_t)c0 * bins;
    for (int o = tid; o < nc * bins; o += kRoiThreads) dst[o] = otile[o];
  }
#undef ROI_STAGE
}

// ------------------------------------------------------------------------------------------
// MultiScaleRoIAlign, register / shuffle formulation (the common case: out * sampling_ratio <= 32 per axis, i.e. the
// reference's 7 x 7 and 14 x 14 poolers with sampling_ratio 2).
//
// The staged kernel above spends its time in the shared-memory pipe (22 M wavefronts per 1000 RoIs: 4-byte cp.async
// writes of the gather, tap and table reads) behind CTA barriers.  Here nothing is staged.  roi_align is separable:
//     bin(ph, pw) = 1/count * sum_{iy, ix} (hy*v[ylo] + ly*v[yhi]) (x-interpolated)
// so a warp takes one channel at a time with LANE = DISTINCT FEATURE COLUMN the RoI touches (at most 2 * out * sr
// of them, in one or two registers per lane): for an output row ph it loads the 2 * sr feature rows its samples need
// straight from global memory (one coalesced <= 128-byte load per row and lane register; rows shared by neighbouring
// samples hit L1), blends them with the y weights into one register T per lane, and then the lanes switch roles --
// lane = x sample (pw, ix) -- and fetch their two taps of T with two shuffles, weight them, add the sr samples of a bin
// with shuffles, and store the bin.  Per (RoI, channel) that is ~2 * out * sr loads and ~20 instructions per output
// row, no shared-memory traffic beyond one broadcast 16-byte read per sample row, and no barrier after the prologue.
// Arithmetic: y blend first, then x, with fused multiply-adds -- a different association than torchvision's
// w1*v1 + w2*v2 + w3*v3 + w4*v4 (well inside the 1e-5 bar; RoIAlign is not an index-producing stage).
// ------------------------------------------------------------------------------------------
constexpr int kRoiRowThreads = 256;
constexpr int kRoiRowWarps = kRoiRowThreads / 32;

struct RowSample { int lo, hi; float h, l; };  // feature-row OFFSETS (row * fw) of a y sample's two taps and their weights

// x phase of one output row: lane = x sample; its two taps of T come from the lanes that own those columns
template <int SR>
__device__ __forceinline__ float roi_row_x_phase(float t0, float t1, bool wide, int plo, int phi, float hx, float lx, int sr) {
  float vlo = __shfl_sync(0xffffffffu, t0, plo & 31), vhi = __shfl_sync(0xffffffffu, t0, phi & 31);
  if (wide) {
    const float wlo = __shfl_sync(0xffffffffu, t1, plo & 31), whi = __shfl_sync(0xffffffffu, t1, phi & 31);
    if (plo >= 32) vlo = wlo;
    if (phi >= 32) vhi = whi;
  }
  float v = fmaf(hx, vlo, lx * vhi);
  if (SR == 2) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
  } else {
    const float own = v;
    for (int k = 1; k < sr; ++k) v += __shfl_down_sync(0xffffffffu, own, k);
  }
  return v;
}

// OH > 0 (with SR > 0): output height known at compile time -- all 2 * OH * SR row loads of a channel are issued before
// anything depends on them (one memory round trip per channel instead of one per output row).
template <int SR, int OH>
__device__ __forceinline__ void roi_rows_body(const RoiLevels& L, const float4* __restrict__ boxes,
                                              const int32_t* __restrict__ batch_idx, int channels, int oh, int ow, int sr_rt,
                                              int cpc, float* __restrict__ out) {
  const int sr = SR > 0 ? SR : sr_rt;
  __shared__ RowSample s_y[32];
  __shared__ int s_xlist[64];          // distinct feature columns, ascending
  __shared__ int s_xpos[32];           // per x sample: position of its lo tap | hi tap << 8 in s_xlist
  __shared__ float s_xl[32];           // per x sample: weight of the hi tap (lx); < 0 marks a sample outside the map
  __shared__ int s_nx;
  const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float4 b = boxes[r];
  int lvl = 0;
  if (L.num_levels > 1) {  // LevelMapper (poolers.py:73-84), every thread: a handful of instructions
    const float s = sqrtf((b.z - b.x) * (b.w - b.y));
    float t = floorf(L.canonical_level + log2f(s / L.canonical_scale) + L.eps);
    t = fminf(fmaxf(t, (float)L.k_min), (float)L.k_max);
    lvl = (int)(long long)t - L.k_min;
  }
  const int fh = L.h[lvl], fw = L.w[lvl];
  const float ss = L.scale[lvl];
  const float start_w = b.x * ss, start_h = b.y * ss, end_w = b.z * ss, end_h = b.w * ss;
  const float roi_w = fmaxf(end_w - start_w, 1.0f), roi_h = fmaxf(end_h - start_h, 1.0f);
  const float bin_h = roi_h / (float)oh, bin_w = roi_w / (float)ow;
  const int ny = oh * sr, nx = ow * sr;
  if (warp == 0 && lane < ny) {
    const int ph = lane / sr, iy = lane - ph * sr;
    const AxisSample a = axis_sample(start_h + (float)ph * bin_h + (((float)iy + 0.5f) * bin_h) / (float)sr, fh);
    RowSample rs;
    rs.lo = a.bad ? 0 : a.lo * fw;
    rs.hi = a.bad ? 0 : a.hi * fw;
    rs.h = a.bad ? 0.0f : a.h;
    rs.l = a.bad ? 0.0f : a.l;
    s_y[lane] = rs;
  }
  if (warp == 1) {
    // x samples are monotone along the axis and a sample's hi tap is its lo tap or the next column, so a tap is new
    // iff it exceeds the previous valid sample's hi tap: positions in the distinct list come from one warp prefix sum
    AxisSample a;
    a.bad = 1; a.lo = a.hi = 0; a.l = a.h = 0.0f;
    if (lane < nx) {
      const int pw = lane / sr, ix = lane - pw * sr;
      a = axis_sample(start_w + (float)pw * bin_w + (((float)ix + 0.5f) * bin_w) / (float)sr, fw);
    }
    const bool ok = !a.bad;
    const unsigned okmask = __ballot_sync(0xffffffffu, ok);
    const int prev_hi = __shfl_up_sync(0xffffffffu, a.hi, 1);
    const bool first = ok && !((okmask >> (lane > 0 ? lane - 1 : 0)) & 1u) ? true : (ok && lane == 0);
    // (valid samples are one contiguous run: the ones outside the map sit at the two ends)
    const bool new_lo = ok && (first || a.lo > prev_hi);
    const bool new_hi = ok && a.hi > a.lo && (first || a.hi > prev_hi);
    const int add = (new_lo ? 1 : 0) + (new_hi ? 1 : 0);
    int incl = add;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    const int base = incl - add;
    int pos_lo = 0, pos_hi = 0;
    if (ok) {
      pos_lo = new_lo ? base : (a.lo == prev_hi ? base - 1 : base - 2);
      pos_hi = a.hi == a.lo ? pos_lo : (new_hi ? base + (new_lo ? 1 : 0) : base - 1);
      if (new_lo) s_xlist[base] = a.lo;
      if (new_hi) s_xlist[base + (new_lo ? 1 : 0)] = a.hi;
    }
    if (lane < 32) {
      s_xpos[lane] = pos_lo | (pos_hi << 8);
      s_xl[lane] = ok ? a.l : -1.0f;
    }
    if (lane == 31) s_nx = incl;
  }
  __syncthreads();
  const int nxl = s_nx;
  // lane as a feature column (two registers cover up to 64 distinct columns)
  // lanes past the last distinct column repeat it (a load of column 0 would pull an unrelated sector into every request)
  const int col0 = nxl > 0 ? s_xlist[min(lane, nxl - 1)] : 0;
  const int col1 = nxl > 32 ? s_xlist[min(lane + 32, nxl - 1)] : col0;
  const bool wide = nxl > 32;
  // lane as an x sample
  const int xp = s_xpos[lane];
  const int plo = xp & 0xff, phi = (xp >> 8) & 0xff;
  const float xl_raw = s_xl[lane];
  const float lx = (lane < nx && xl_raw >= 0.0f) ? xl_raw : 0.0f;
  const float hx = (lane < nx && xl_raw >= 0.0f) ? 1.0f - xl_raw : 0.0f;
  const int my_pw = lane / sr;
  const bool bin_lane = lane < nx && lane - my_pw * sr == 0;
  const float inv_count = 1.0f / (float)(sr * sr);
  const int bins = oh * ow;
  const int img = batch_idx ? batch_idx[r] : 0;
  const size_t fhw = (size_t)fh * fw;
  const float* fbase = L.feat[lvl] + (size_t)img * channels * fhw;
  float* obase = out + (size_t)r * channels * bins;
  const int cbase = blockIdx.y * cpc, cend = min(channels, cbase + cpc);
  for (int c = cbase + warp; c < cend; c += kRoiRowWarps) {
    const float* pl = fbase + (size_t)c * fhw;
    float* oc = obase + (size_t)c * bins;
    if constexpr (OH > 0 && SR > 0) {
      // output rows in chunks of CH, TWO channels per pass: 4 * CH * SR independent row loads in flight per lane (the
      // kernel is bound by memory latency x sequential passes per warp, not by bytes: more loads per pass, fewer passes),
      // then the chunk's y blend and x phase.  The chunk loop is not unrolled, so a 14-row pooler needs no more registers
      // than a 7-row one; the sample table is re-read from shared memory (broadcast) instead of living in registers.
      constexpr int CH = OH % 7 == 0 ? 7 : OH;
      constexpr int NYC = CH * SR;
      const bool two = c + kRoiRowWarps < cend;            // second channel of the pass (warp-uniform)
      const float* pl2 = two ? pl + (size_t)kRoiRowWarps * fhw : pl;
#pragma unroll 1
      for (int ph0 = 0; ph0 < OH; ph0 += CH) {
        const RowSample* sy = s_y + ph0 * SR;
#pragma unroll 1
        for (int half = 0; half < (wide ? 2 : 1); ++half) {
          const int col = half ? col1 : col0;
          float va[2 * NYC], vb[2 * NYC];
#pragma unroll
          for (int q = 0; q < NYC; ++q) {
            const int lo = sy[q].lo + col, hi = sy[q].hi + col;
            va[2 * q] = __ldg(pl + lo);
            va[2 * q + 1] = __ldg(pl + hi);
            vb[2 * q] = __ldg(pl2 + lo);
            vb[2 * q + 1] = __ldg(pl2 + hi);
          }
          // y blend; with `wide` the x phase needs both column halves of a row: the first half's rows are parked in
          // shared-memory-free fashion by running the x phase per half and adding (the x weights select by position)
#pragma unroll
          for (int ph = 0; ph < CH; ++ph) {
            float ta = 0.0f, tb = 0.0f;
#pragma unroll
            for (int iy = 0; iy < SR; ++iy) {
              const float wh = sy[ph * SR + iy].h, wl = sy[ph * SR + iy].l;
              ta = fmaf(wh, va[2 * (ph * SR + iy)], ta);
              ta = fmaf(wl, va[2 * (ph * SR + iy) + 1], ta);
              tb = fmaf(wh, vb[2 * (ph * SR + iy)], tb);
              tb = fmaf(wl, vb[2 * (ph * SR + iy) + 1], tb);
            }
            // x phase restricted to the taps that live in this column half
            const bool lo_here = (plo >= 32) == (half == 1), hi_here = (phi >= 32) == (half == 1);
            const float alo = __shfl_sync(0xffffffffu, ta, plo & 31), ahi = __shfl_sync(0xffffffffu, ta, phi & 31);
            const float blo = __shfl_sync(0xffffffffu, tb, plo & 31), bhi = __shfl_sync(0xffffffffu, tb, phi & 31);
            float xa = (lo_here ? hx * alo : 0.0f) + (hi_here ? lx * ahi : 0.0f);
            float xb = (lo_here ? hx * blo : 0.0f) + (hi_here ? lx * bhi : 0.0f);
            if (SR == 2) {
              xa += __shfl_xor_sync(0xffffffffu, xa, 1);
              xb += __shfl_xor_sync(0xffffffffu, xb, 1);
            } else {
              const float oa = xa, ob = xb;
              for (int k = 1; k < sr; ++k) {
                xa += __shfl_down_sync(0xffffffffu, oa, k);
                xb += __shfl_down_sync(0xffffffffu, ob, k);
              }
            }
            if (bin_lane) {
              float* o1 = oc + (ph0 + ph) * ow + my_pw;
              float* o2 = o1 + (size_t)kRoiRowWarps * bins;
              if (half == 0) {
                *o1 = xa * inv_count;
                if (two) *o2 = xb * inv_count;
              } else {  // second column half of a wide RoI: add to what the first half stored
                *o1 += xa * inv_count;
                if (two) *o2 += xb * inv_count;
              }
            }
          }
        }
      }
      c += kRoiRowWarps;  // the pass covered channels c and c + kRoiRowWarps
      continue;
    }
#pragma unroll 2
    for (int ph = 0; ph < oh; ++ph) {
      float t0 = 0.0f, t1 = 0.0f;
#pragma unroll
      for (int iy = 0; iy < (SR > 0 ? SR : 1); ++iy) {
        const RowSample rs = s_y[ph * sr + iy];
        const float a = __ldg(pl + rs.lo + col0), bb = __ldg(pl + rs.hi + col0);
        t0 = fmaf(rs.h, a, t0);
        t0 = fmaf(rs.l, bb, t0);
        if (wide) {
          const float a1 = __ldg(pl + rs.lo + col1), b1 = __ldg(pl + rs.hi + col1);
          t1 = fmaf(rs.h, a1, t1);
          t1 = fmaf(rs.l, b1, t1);
        }
      }
      if (SR == 0) {
        for (int iy = 1; iy < sr; ++iy) {
          const RowSample rs = s_y[ph * sr + iy];
          t0 = fmaf(rs.h, __ldg(pl + rs.lo + col0), t0);
          t0 = fmaf(rs.l, __ldg(pl + rs.hi + col0), t0);
          if (wide) {
            t1 = fmaf(rs.h, __ldg(pl + rs.lo + col1), t1);
            t1 = fmaf(rs.l, __ldg(pl + rs.hi + col1), t1);
          }
        }
      }
      const float v = roi_row_x_phase<SR>(t0, t1, wide, plo, phi, hx, lx, sr);
      if (bin_lane) oc[ph * ow + my_pw] = v * inv_count;
    }
  }
}

template <int SR, int OH>
__global__ void __launch_bounds__(kRoiRowThreads, 2) ms_roi_align_rows_kernel(RoiLevels L, const float4* __restrict__ boxes,
                                                                        const int32_t* __restrict__ batch_idx, int channels,
                                                                        int oh, int ow, int sr_rt, int cpc,
                                                                        float* __restrict__ out) {
  roi_rows_body<SR, OH>(L, boxes, batch_idx, channels, oh, ow, sr_rt, cpc, out);
}

// ------------------------------------------------------------------------------------------
// TMA-staged RoIAlign (sampling ratio 2, 7 x 7 and 14 x 14 poolers).  What a RoI needs of one channel is a WINDOW of its
// FPN level -- the rows and columns between its first and last bilinear tap, at most ~30 x 30 because the LevelMapper
// picks the level on which the RoI is 14-28 feature pixels across.  Here that window arrives as ONE tensor-tile TMA
// load per (RoI, channel) (cp.async.bulk.tensor.3d, SASS UTMALDG): a tensor map per level and per box shape
// (8/16/24/32 columns x 8/16/24/32 rows x 1 channel, starting at a 16-byte-aligned column; the smallest shape that
// covers the window is used, so small RoIs
// do not drag 4 KB tiles through the L2), zero-filled outside the map.  Every WARP runs its own three-deep ring of
// tiles on its own mbarriers (no CTA barrier after the prologue): lane 0 issues the loads of the warp's next channels
// while the warp reduces the current one out of shared memory with the arithmetic of the rows kernel above (lane =
// window column for the y blend, lane = x sample for the x phase; identical bits).  No load instruction, address
// arithmetic or scoreboard wait for global memory is left in the SM's instruction stream: ~125 instead of ~290 warp
// instructions per (RoI, channel).  RoIs whose window does not fit 32 x 32 (elongated boxes) and levels whose row pitch
// is not a multiple of 16 bytes (42 floats at the coarsest level of an 800 x 1344 image) take roi_rows_body.
// ------------------------------------------------------------------------------------------
constexpr int kTmaStages = 3;
constexpr int kTmaTileBytes = 32 * 32 * 4;
constexpr int kTmaLevels = 4;

struct RoiTmaMaps {
  CUtensorMap m[kTmaLevels][16];  // [level][(rows / 8 - 1) * 4 + (cols / 8 - 1)]
  int ok[8];                      // level usable (pitch and base 16-byte aligned, map encoded)
};

// Planes (channels) per tile.  Several planes per request -- as many as fit a 4 KB ring slot, up to 8 for 8 x 8 windows --
// were measured and change nothing (97.6 vs 95.6 us for 1000 small RoIs; 14 x 14: 69 vs 54 us, the plane offset costs
// three address instructions per tap): the copy unit's request rate is not what bounds the kernel, so one plane it is.
__host__ __device__ constexpr int roi_tma_planes(int, int) { return 1; }

__device__ __forceinline__ void tma_load_3d(void* dst_smem, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
          smem_u32(dst_smem)),
      "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}

template <int OH>
__global__ void __launch_bounds__(kRoiRowThreads, 2)
ms_roi_align_tma_kernel(const __grid_constant__ RoiTmaMaps M, RoiLevels L, const float4* __restrict__ boxes,
                        const int32_t* __restrict__ batch_idx, int channels, int cpc, float* __restrict__ out) {
  constexpr int SR = 2, NS = OH * SR;  // samples per axis (14 or 28)
  extern __shared__ __align__(128) unsigned char tma_smem[];  // [warps][stages][4 KB]
  __shared__ __align__(8) uint64_t bars[kRoiRowWarps][kTmaStages];   // tile landed (TMA transaction bytes)
  __shared__ __align__(8) uint64_t empty[kRoiRowWarps][kTmaStages];  // tile consumed (all 32 lanes arrived)
  __shared__ int s_ylo[32], s_yhi[32], s_xlo[32], s_xhi[32];  // absolute tap rows / columns
  __shared__ float s_yh[32], s_yl[32], s_xl[32];              // weights (0 for samples outside the map); s_xl < 0: outside
  __shared__ int s_win[4];                                    // first / last tap column, first / last tap row
  const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float4 b = boxes[r];
  int lvl = 0;
  if (L.num_levels > 1) {  // LevelMapper (poolers.py:73-84)
    const float s = sqrtf((b.z - b.x) * (b.w - b.y));
    float t = floorf(L.canonical_level + log2f(s / L.canonical_scale) + L.eps);
    t = fminf(fmaxf(t, (float)L.k_min), (float)L.k_max);
    lvl = (int)(long long)t - L.k_min;
  }
  const int fh = L.h[lvl], fw = L.w[lvl];
  const float ss = L.scale[lvl];
  const float start_w = b.x * ss, start_h = b.y * ss, end_w = b.z * ss, end_h = b.w * ss;
  const float roi_w = fmaxf(end_w - start_w, 1.0f), roi_h = fmaxf(end_h - start_h, 1.0f);
  const float bin_h = roi_h / (float)OH, bin_w = roi_w / (float)OH;
  if (tid < kRoiRowWarps * kTmaStages) {
    mbar_init(&bars[tid / kTmaStages][tid % kTmaStages], 1);
    mbar_init(&empty[tid / kTmaStages][tid % kTmaStages], 32);
  }
  if (warp < 2) {  // warp 0: y samples, warp 1: x samples
    const bool isx = warp == 1;
    AxisSample a;
    a.bad = 1; a.lo = a.hi = 0; a.l = a.h = 0.0f;
    if (lane < NS) {
      const int pq = lane / SR, iq = lane - pq * SR;
      a = isx ? axis_sample(start_w + (float)pq * bin_w + (((float)iq + 0.5f) * bin_w) / (float)SR, fw)
              : axis_sample(start_h + (float)pq * bin_h + (((float)iq + 0.5f) * bin_h) / (float)SR, fh);
    }
    const int first = __reduce_min_sync(0xffffffffu, a.bad ? 0x7fffffff : a.lo);
    const int last = __reduce_max_sync(0xffffffffu, a.bad ? -1 : a.hi);
    if (isx) {
      s_xlo[lane] = a.bad ? 0 : a.lo;
      s_xhi[lane] = a.bad ? 0 : a.hi;
      s_xl[lane] = a.bad ? -1.0f : a.l;
      if (lane == 0) { s_win[0] = first; s_win[1] = last; }
    } else {
      s_ylo[lane] = a.bad ? 0 : a.lo;
      s_yhi[lane] = a.bad ? 0 : a.hi;
      s_yh[lane] = a.bad ? 0.0f : a.h;
      s_yl[lane] = a.bad ? 0.0f : a.l;
      if (lane == 0) { s_win[2] = first; s_win[3] = last; }
    }
  }
  __syncthreads();
  // the innermost tile coordinate must be a multiple of 16 bytes (an unaligned one faults with "illegal instruction":
  // tools/tma_probe2.cu): the window starts at the 4-column boundary at or below the first tap
  const int x0 = s_win[0] & ~3, x1 = s_win[1], y0 = s_win[2], y1 = s_win[3];
  const int spanx = x1 - x0 + 1, spany = y1 - y0 + 1;
  const bool fits = lvl < kTmaLevels && M.ok[lvl] && x1 >= x0 && y1 >= y0 && spanx <= 32 && spany <= 32;
  if (!fits) {  // CTA-uniform
    roi_rows_body<SR, OH>(L, boxes, batch_idx, channels, OH, OH, SR, cpc, out);
    return;
  }
  const int bw = (spanx + 7) & ~7, bh = (spany + 7) & ~7;
  constexpr int bc = roi_tma_planes(0, 0);  // planes per tile (see roi_tma_planes)
  if (channels < bc) {  // (a tile must not name planes past the image's last channel)
    roi_rows_body<SR, OH>(L, boxes, batch_idx, channels, OH, OH, SR, cpc, out);
    return;
  }
  const CUtensorMap* map = &M.m[lvl][((bh >> 3) - 1) * 4 + ((bw >> 3) - 1)];
  const int plane = bw * bh;  // floats
  const uint32_t tile_bytes = (uint32_t)(plane * 4 * bc);
  // lane as a window column (y blend) and as an x sample (x phase)
  const int lc = min(lane, bw - 1);
  const int plo = (lane < NS ? s_xlo[lane] - x0 : 0) & 31, phi = (lane < NS ? s_xhi[lane] - x0 : 0) & 31;
  const float xl_raw = lane < NS ? s_xl[lane] : -1.0f;
  const float lx = xl_raw >= 0.0f ? xl_raw : 0.0f;
  const float hx = xl_raw >= 0.0f ? 1.0f - xl_raw : 0.0f;
  const int my_pw = lane / SR;
  const bool bin_lane = lane < NS && lane - my_pw * SR == 0;
  const float inv_count = 1.0f / (float)(SR * SR);
  constexpr int bins = OH * OH;
  const int img = batch_idx ? batch_idx[r] : 0;
  float* obase = out + (size_t)r * channels * bins;
  const int cbase = blockIdx.y * cpc, cend = min(channels, cbase + cpc);
  // this warp's tiles: blocks of bc channels, block index warp, warp + 8, ...
  const int nblk = (cend - cbase + bc - 1) / bc;
  const int nk = warp < nblk ? (nblk - warp + kRoiRowWarps - 1) / kRoiRowWarps : 0;
  // the warp's ring: kTmaStages slots of 4 KB.  The tile loop is unrolled by the ring depth, so a tap's shared-memory
  // address is a per-RoI register plus an IMMEDIATE stage offset (plus the plane offset inside a multi-channel tile)
  unsigned char* wbuf = tma_smem + (size_t)warp * kTmaStages * kTmaTileBytes;
  uint64_t* wbar = bars[warp];
  uint64_t* wempty = empty[warp];
  auto first_plane = [&](int k) {  // first channel of tile k, pulled back so that the tile ends inside the image's planes
    return min(cbase + (warp + k * kRoiRowWarps) * bc, channels - bc);
  };
  auto issue = [&](int k, int st) {  // lane 0: tile k of this warp -> stage st
    mbar_expect_tx(&wbar[st], tile_bytes);
    tma_load_3d(wbuf + (size_t)st * kTmaTileBytes, map, x0, y0, img * channels + first_plane(k), &wbar[st]);
  };
  if (lane == 0)
    for (int k = 0; k < kTmaStages - 1 && k < nk; ++k) issue(k, k);
  // y taps: addresses of this lane's column in stage 0, plane 0, with their weights; the x weights carry 1 / count (a power of two)
  const float* T0 = reinterpret_cast<const float*>(wbuf) + lc;
  const float* alo_p[OH == 7 ? NS : 1];
  const float* ahi_p[OH == 7 ? NS : 1];
  float wh[OH == 7 ? NS : 1], wl[OH == 7 ? NS : 1];
  if constexpr (OH == 7) {
#pragma unroll
    for (int q = 0; q < NS; ++q) {
      alo_p[q] = T0 + (s_ylo[q] - y0) * bw;
      ahi_p[q] = T0 + (s_yhi[q] - y0) * bw;
      wh[q] = s_yh[q];
      wl[q] = s_yl[q];
    }
  }
  const float hxc = hx * inv_count, lxc = lx * inv_count;
  uint32_t par = 0;
  for (int k0 = 0; k0 < nk; k0 += kTmaStages) {
#pragma unroll
    for (int st = 0; st < kTmaStages; ++st) {
      const int k = k0 + st;
      if (k < nk) {  // warp-uniform
        // the load kTmaStages - 1 tiles ahead goes into the stage the previous tile released: its 32 lanes arrived on the
        // stage's "empty" barrier after their last read -- the release / acquire pair orders those generic-proxy reads
        // before the async-proxy refill (no fence.proxy.async, which lowers to MEMBAR.ALL.CTA and waits for the stores)
        if (lane == 0 && k + kTmaStages - 1 < nk) {
          const int s2 = (st + kTmaStages - 1) % kTmaStages;
          if (st == 0) { if (k0 > 0) mbar_wait(&wempty[s2], par ^ 1u); } else mbar_wait(&wempty[s2], par);
          issue(k + kTmaStages - 1, s2);
        }
        mbar_wait(&wbar[st], par);
        constexpr int kStageFloats = kTmaTileBytes / 4;
        const int c_first = first_plane(k);                                  // plane 0 of the tile
        const int c_lo = cbase + (warp + k * kRoiRowWarps) * bc;              // channels this tile is responsible for
        const int c_hi = min(c_lo + bc, cend);
#pragma unroll 1
        for (int c = c_lo; c < c_hi; ++c) {
          const int poff = (c - c_first) * plane;
          float* oc = obase + (size_t)c * bins;
#pragma unroll
          for (int ph = 0; ph < OH; ++ph) {
            float ta = 0.0f;
#pragma unroll
            for (int iy = 0; iy < SR; ++iy) {
              const int q = ph * SR + iy;
              float vlo, vhi, h_, l_;
              if constexpr (OH == 7) {
                vlo = (alo_p[q] + poff)[st * kStageFloats]; vhi = (ahi_p[q] + poff)[st * kStageFloats]; h_ = wh[q]; l_ = wl[q];
              } else {
                vlo = T0[st * kStageFloats + poff + (s_ylo[q] - y0) * bw]; vhi = T0[st * kStageFloats + poff + (s_yhi[q] - y0) * bw];
                h_ = s_yh[q]; l_ = s_yl[q];
              }
              ta = fmaf(h_, vlo, ta);
              ta = fmaf(l_, vhi, ta);
            }
            const float alo = __shfl_sync(0xffffffffu, ta, plo), ahi = __shfl_sync(0xffffffffu, ta, phi);
            float xa = fmaf(hxc, alo, lxc * ahi);
            xa += __shfl_xor_sync(0xffffffffu, xa, 1);
            if (bin_lane) oc[ph * OH + my_pw] = xa;
          }
        }
        mbar_arrive(&wempty[st]);  // this lane's reads of the stage are done (release)
      }
    }
    par ^= 1u;
  }
}

// ------------------------------------------------------------------------------------------
// GeneralizedRCNNTransform for inference, one image (transform.py:436-459): normalize (:264-270, float32
// (v - mean) / std), resize (:284-315 -> F.interpolate bilinear, align_corners=False, output size given) and the
// image's slot of batch_images (:352-386: zero canvas, image at the top-left) in one pass: a thread per canvas
// pixel of a channel plane reads its four taps, normalises each (the reference interpolates normalised values) and
// writes the pixel, or writes 0 outside the resized image.  out = h0*(w0*p00 + w1*p01) + h1*(w0*p10 + w1*p11).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rcnn_transform_kernel(const float* __restrict__ img, int in_h, int in_w, int out_h, int out_w, float m0, float m1, float m2,
                      float s0, float s1, float s2, float* __restrict__ canvas, int can_h, int can_w) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, c = blockIdx.z;
  if (x >= can_w) return;
  float r = 0.0f;
  if (y < out_h && x < out_w) {
    const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2), sd = c == 0 ? s0 : (c == 1 ? s1 : s2);
    const float ry = (float)in_h / (float)out_h, rx = (float)in_w / (float)out_w;
    int iy0, iy1, ix0, ix1;
    float ly0, ly1, lx0, lx1;
    bilinear_src(y, ry, in_h, iy0, iy1, ly0, ly1);
    bilinear_src(x, rx, in_w, ix0, ix1, lx0, lx1);
    const float* pl = img + (size_t)c * in_h * in_w;
    const float p00 = (__ldg(pl + (size_t)iy0 * in_w + ix0) - mean) / sd, p01 = (__ldg(pl + (size_t)iy0 * in_w + ix1) - mean) / sd;
    const float p10 = (__ldg(pl + (size_t)iy1 * in_w + ix0) - mean) / sd, p11 = (__ldg(pl + (size_t)iy1 * in_w + ix1) - mean) / sd;
    r = ly0 * (lx0 * p00 + lx1 * p01) + ly1 * (lx0 * p10 + lx1 * p11);
  }
  canvas[((size_t)c * can_h + y) * can_w + x] = r;
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
int configure_rcnn_ops(isb_handle* h) {  // once per handle (isb_create), on the handle's device
  ISB_CUDA(h, cudaFuncSetAttribute(paste_masks_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPasteSmem));
  return ISB_OK;
}
int launch_paste_masks(isb_handle* h, const float* masks, const float* boxes, int m, int S, int pad, int im_h, int im_w,
                       float* out, cudaStream_t s) {
  const int64_t total = (int64_t)m * im_h * im_w;  // `out` is 16-byte aligned (checked by isb_paste_masks)
  if (total <= 0) return ISB_OK;
  const int64_t nchunks = (total + kPasteChunk - 1) / kPasteChunk;
  const int grid = (int)min((int64_t)h->num_sms, (nchunks + kPasteWarps - 1) / kPasteWarps);
  paste_masks_bulk_kernel<<<grid, kPasteWarps * 32, kPasteSmem, s>>>(masks, reinterpret_cast<const float4*>(boxes), m, S, pad, im_h,
                                                                    im_w, out);
  ISB_LAUNCH_CHECK(h, "paste_masks_bulk_kernel");
  return ISB_OK;
}
int launch_paste_class_map(isb_handle* h, const float* masks, const float* boxes, const int64_t* labels, const float* scores, int m,
                           int S, int pad, int paste_h, int paste_w, int out_h, int out_w, float confidence, uint8_t* class_map,
                           int64_t* masks_arg, int32_t* kept_count, cudaStream_t s) {
  const size_t smem = (size_t)(m > 0 ? m : 1) * sizeof(PasteDet);
  if (smem > 160 * 1024) ISB_FAIL(h, ISB_ERR_LIMIT, "paste_class_map: %d detections exceed shared memory", m);
  if (smem > 48 * 1024)
    ISB_CUDA(h, cudaFuncSetAttribute(paste_class_map_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t total = (int64_t)out_h * out_w;
  paste_class_map_kernel<<<grid_for(total, kPcmThreads, h->num_sms, 8), kPcmThreads, smem, s>>>(
      masks, reinterpret_cast<const float4*>(boxes), labels, scores, m, S, pad, paste_h, paste_w, out_h, out_w, confidence, class_map,
      masks_arg, kept_count);
  ISB_LAUNCH_CHECK(h, "paste_class_map_kernel");
  return ISB_OK;
}
int launch_rcnn_transform(isb_handle* h, const float* img, int in_h, int in_w, int out_h, int out_w, const float mean[3],
                          const float stdv[3], float* canvas, int can_h, int can_w, cudaStream_t s) {
  rcnn_transform_kernel<<<dim3((can_w + 255) / 256, can_h, 3), 256, 0, s>>>(img, in_h, in_w, out_h, out_w, mean[0], mean[1], mean[2],
                                                                            stdv[0], stdv[1], stdv[2], canvas, can_h, can_w);
  ISB_LAUNCH_CHECK(h, "rcnn_transform_kernel");
  return ISB_OK;
}
// Tensor maps of the TMA RoIAlign: one per level and box shape, encoded on the host (cuTensorMapEncodeTiled through the
// runtime's driver entry point: libisb.so does not link libcuda) and kept per device for as long as the feature maps stay
// where they are -- the steady state of a loop over preallocated tensors.  nullptr: no usable level / no driver entry
// point (the caller then runs the rows kernel).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
struct RoiTmaCache {
  RoiTmaMaps maps;
  const float* feat[kTmaLevels];
  int h[kTmaLevels], w[kTmaLevels], levels, channels, valid, any;
};
// Copies the maps into *out under the cache lock (another thread's call on the same device may re-encode the entry).
static bool roi_tma_maps(isb_handle* h, const RoiLevels& L, int channels, RoiTmaMaps* out) {
  static std::mutex mu;
  static RoiTmaCache cache[64];
  static EncodeTiledFn encode = nullptr;
  static bool looked = false;
  std::lock_guard<std::mutex> lock(mu);
  if (!looked) {
    looked = true;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      encode = reinterpret_cast<EncodeTiledFn>(fn);
    cudaGetLastError();
  }
  if (!encode || h->device < 0 || h->device >= 64) return false;
  RoiTmaCache& c = cache[h->device];
  const int nl = L.num_levels < kTmaLevels ? L.num_levels : kTmaLevels;
  bool same = c.valid && c.levels == nl && c.channels == channels;
  for (int l = 0; same && l < nl; ++l) same = c.feat[l] == L.feat[l] && c.h[l] == L.h[l] && c.w[l] == L.w[l];
  if (same) {
    if (c.any) *out = c.maps;
    return c.any != 0;
  }
  memset(&c, 0, sizeof(c));
  c.levels = nl;
  c.channels = channels;
  for (int l = 0; l < nl; ++l) {
    c.feat[l] = L.feat[l];
    c.h[l] = L.h[l];
    c.w[l] = L.w[l];
    if ((L.w[l] & 3) || (reinterpret_cast<uintptr_t>(L.feat[l]) & 15) || L.w[l] < 1 || L.h[l] < 1) continue;
    // [planes = images x channels][rows][columns]; the plane count is only an out-of-bounds limit (the image count is not
    // an argument of the call: any plane the kernel names exists)
    const cuuint64_t dims[3] = {(cuuint64_t)L.w[l], (cuuint64_t)L.h[l], (cuuint64_t)channels * 65536ull};
    const cuuint64_t strides[2] = {(cuuint64_t)L.w[l] * 4ull, (cuuint64_t)L.w[l] * (cuuint64_t)L.h[l] * 4ull};
    const cuuint32_t estr[3] = {1, 1, 1};
    bool ok = true;
    for (int by = 0; ok && by < 4; ++by)
      for (int bx = 0; ok && bx < 4; ++bx) {
        const cuuint32_t box[3] = {(cuuint32_t)(8 * (bx + 1)), (cuuint32_t)(8 * (by + 1)), (cuuint32_t)roi_tma_planes(8 * (bx + 1), 8 * (by + 1))};
        ok = encode(&c.maps.m[l][by * 4 + bx], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(L.feat[l]), dims, strides, box,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
      }
    c.maps.ok[l] = ok ? 1 : 0;
    c.any |= c.maps.ok[l];
  }
  c.valid = 1;
  if (c.any) *out = c.maps;
  return c.any != 0;
}

int launch_ms_roi_align(isb_handle* h, const RoiLevels& L, const float* boxes, const int32_t* batch_idx, int num_rois,
                        int channels, int oh, int ow, int sr, float* out, cudaStream_t s) {
  if (!h->roi_configured) {
    ISB_CUDA(h, cudaFuncSetAttribute(ms_roi_align_tma_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kRoiRowWarps * kTmaStages * kTmaTileBytes));
    ISB_CUDA(h, cudaFuncSetAttribute(ms_roi_align_tma_kernel<14>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kRoiRowWarps * kTmaStages * kTmaTileBytes));
    const int max_smem = (int)(kRoiTileFloats * sizeof(float) + (size_t)kRoiMaxEntries * 32);
    ISB_CUDA(h, cudaFuncSetAttribute(ms_roi_align_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
    ISB_CUDA(h, cudaFuncSetAttribute(ms_roi_align_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
    h->roi_configured = 1;
  }
  static const bool force_staged = getenv("ISB_ROI_STAGED") != nullptr;  // A/B switch: the round-1 staged kernel
  if (oh * sr <= 32 && ow * sr <= 32 && !force_staged) {
    // register / shuffle kernel: a CTA = (RoI, group of channels), a warp walks the group's channels.  Enough CTAs for
    // ~2 waves of 8 resident CTAs per SM; at least one channel per warp and pass.
    // (swept for the TMA kernel at 1000 RoIs: one group of 256 channels per RoI 113 us, two or three groups 96 us)
    int groups = (16 * h->num_sms + num_rois - 1) / num_rois;
    if (groups > (channels + kRoiRowWarps - 1) / kRoiRowWarps) groups = (channels + kRoiRowWarps - 1) / kRoiRowWarps;
    if (groups < 1) groups = 1;
    int cpc = (channels + groups - 1) / groups;
    cpc = (cpc + 2 * kRoiRowWarps - 1) / (2 * kRoiRowWarps) * (2 * kRoiRowWarps);  // a warp takes channels in pairs
    groups = (channels + cpc - 1) / cpc;
    const float4* b4 = reinterpret_cast<const float4*>(boxes);
    static const bool no_tma = getenv("ISB_ROI_NO_TMA") != nullptr;  // A/B switch: the register / shuffle rows kernel alone
    // the 7 x 7 box-head pooler: faster than the rows kernel on every box population measured (101 / 104 / 122 us against
    // 123 / 124 / 127, tools/bench_roi_shapes.py).  The 14 x 14 mask-head pooler stays on the rows kernel unless
    // ISB_ROI_TMA14 is set: its 28-sample table does not fit in registers next to the tile pointers and is re-read from
    // shared memory per tap (54 us against 47.5 us for 100 RoIs; 33.5 against 36.3 on the chain bench's detections).
    static const bool tma14 = getenv("ISB_ROI_TMA14") != nullptr;
    if (sr == 2 && oh == ow && (oh == 7 || (oh == 14 && tma14)) && !no_tma) {
      RoiTmaMaps maps_copy;
      const RoiTmaMaps* maps = &maps_copy;
      if (roi_tma_maps(h, L, channels, &maps_copy)) {
        const size_t smem = (size_t)kRoiRowWarps * kTmaStages * kTmaTileBytes;
        if (oh == 7)
          ms_roi_align_tma_kernel<7><<<dim3(num_rois, groups), kRoiRowThreads, smem, s>>>(*maps, L, b4, batch_idx, channels, cpc, out);
        else
          ms_roi_align_tma_kernel<14><<<dim3(num_rois, groups), kRoiRowThreads, smem, s>>>(*maps, L, b4, batch_idx, channels, cpc, out);
        ISB_LAUNCH_CHECK(h, "ms_roi_align_tma_kernel");
        return ISB_OK;
      }
    }
    if (sr == 2 && oh == 7)        // box head pooler (faster_rcnn_framework.py:326)
      ms_roi_align_rows_kernel<2, 7><<<dim3(num_rois, groups), kRoiRowThreads, 0, s>>>(L, b4, batch_idx, channels, oh, ow, sr, cpc, out);
    else if (sr == 2 && oh == 14)  // mask head pooler (mask_rcnn.py:150)
      ms_roi_align_rows_kernel<2, 14><<<dim3(num_rois, groups), kRoiRowThreads, 0, s>>>(L, b4, batch_idx, channels, oh, ow, sr, cpc, out);
    else if (sr == 2)
      ms_roi_align_rows_kernel<2, 0><<<dim3(num_rois, groups), kRoiRowThreads, 0, s>>>(L, b4, batch_idx, channels, oh, ow, sr, cpc, out);
    else
      ms_roi_align_rows_kernel<0, 0><<<dim3(num_rois, groups), kRoiRowThreads, 0, s>>>(L, b4, batch_idx, channels, oh, ow, sr, cpc, out);
    ISB_LAUNCH_CHECK(h, "ms_roi_align_rows_kernel");
    return ISB_OK;
  }
  // larger out * sampling_ratio: the staged kernel.
  // shared memory: the tile budget + a (bin, sample) table of exactly this call's size (more entries than the cap: the
  // kernel's direct-gather path, no table)
  const int tile_floats = kRoiTileFloats;
  int table_cap = oh * ow * sr * sr;
  if (table_cap > kRoiMaxEntries) table_cap = 0;
  const size_t smem = (size_t)tile_floats * sizeof(float) + (size_t)table_cap * 32;
  // channel groups per RoI: enough CTAs to fill the GPU about a dozen times over, whole passes of 32 channels each
  int groups = (12 * h->num_sms + num_rois - 1) / num_rois;
  if (groups > (channels + 31) / 32) groups = (channels + 31) / 32;
  if (groups < 1) groups = 1;
  int cpc = (channels + groups - 1) / groups;
  cpc = (cpc + 31) & ~31;
  groups = (channels + cpc - 1) / cpc;
  if (sr == 2)
    ms_roi_align_kernel<2><<<dim3(num_rois, groups), kRoiThreads, smem, s>>>(
        L, reinterpret_cast<const float4*>(boxes), batch_idx, channels, oh, ow, sr, cpc, tile_floats, table_cap, out);
  else
    ms_roi_align_kernel<0><<<dim3(num_rois, groups), kRoiThreads, smem, s>>>(
        L, reinterpret_cast<const float4*>(boxes), batch_idx, channels, oh, ow, sr, cpc, tile_floats, table_cap, out);
  ISB_LAUNCH_CHECK(h, "ms_roi_align_kernel");
  return ISB_OK;
}

}  // namespace isb
