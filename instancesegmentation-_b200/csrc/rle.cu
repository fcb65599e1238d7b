// COCO run-length encoding of the instance masks on the device.
//
// The reference hands every instance mask to pycocotools on the host (inst_model/yolact/utils/utils_map.py:33-34:
// `pycocotools.mask.encode(np.asfortranarray(segmentation.astype(np.uint8)))`, driven per detection by prep_metrics
// :56-62), after copying bool [H,W,m] (30 MB at 550^2 x 100) over PCIe.  pycocotools is a third-party dependency that is
// not vendored in the reference (no version pinned anywhere in the tree; the algorithm restated here is maskApi.c
// rleEncode / rleToString of cocoapi 2.0): the mask is scanned in COLUMN-major order and `counts` holds the lengths of
// the alternating runs, starting with a (possibly empty) run of zeros.
//
// Device side: a run starts wherever a pixel differs from its predecessor in column-major order (the predecessor of the
// first pixel is 0).  With the masks laid out [H,W,m] (m innermost) the entries e = x*m + j of one image row are
// contiguous, so thread e walks DOWN its column (x, j) with fully coalesced byte loads across the warp, carrying the
// previous value in a register:
//   A  rle_count_kernel   boundaries per (column, detection)
//   B  rle_scan_kernel    per detection: exclusive prefix over its columns, total
//   B2 rle_offsets_kernel exclusive prefix over detections -> where each detection's counts start in the flat output
//   C  rle_emit_kernel    the same walk again (the image is L2-resident by now), writing boundary positions x*H + y
//   D  rle_diff_kernel    counts = differences of consecutive positions (+ the leading zero run and the final run)
// The counts -> ASCII step (rleToString) is byte-serial string building and stays on the host: isb_rle_to_string.
#include "isb_common.cuh"

namespace isb {

constexpr int kRleThreads = 256;

struct RleParams {
  const uint8_t* masks;      // B slots
  int64_t slot_bytes;
  const int32_t* counts;     // [B] device, or nullptr with m_fixed
  int32_t m_fixed;
  int32_t det_stride;        // D: rows of the per-detection outputs per image
  int32_t batch, hgt, wid;
};

__device__ __forceinline__ int rle_m(const RleParams& p, int b) {
  const int m = p.counts ? p.counts[b] : p.m_fixed;
  return m < 0 ? 0 : min(m, p.det_stride);
}

// A / C share the walk.  EMIT = false: col_cnt[b][e] = number of boundaries in column e.  EMIT = true: col_cnt holds the
// exclusive prefix within the detection; positions go to pos[pos_off[b][j] + prefix + k].
// A thread owns V consecutive columns e (V = 8, 4, 2 or 1: the widest that divides the image's row length W*m, so that
// every row of the thread's word is aligned) and loads them as one V-byte word per image row.
template <int V> struct RleWord;
template <> struct RleWord<1> { using T = uint8_t; };
template <> struct RleWord<2> { using T = uint16_t; };
template <> struct RleWord<4> { using T = uint32_t; };
template <> struct RleWord<8> { using T = uint64_t; };

// Mask bytes are 0 / 1 (bool), so "differs from the pixel above" for all V columns of a word is ONE xor with the
// previous row's word, and the count pass adds those flag bytes lane-wise (flushed to 32-bit counters every 255 rows):
// about three instructions per V pixels, which leaves the walk to the memory system.  Boundaries are sparse (0.06 % of
// the pixels), so the emit pass only looks inside a word when its xor is non-zero.
template <bool EMIT, int V>
__device__ __forceinline__ void rle_walk(const RleParams& p, int b, int m, int e0, int32_t* __restrict__ col_cnt,
                                         const int64_t* __restrict__ pos_off, uint32_t* __restrict__ pos, int64_t cap) {
  using W = typename RleWord<V>::T;
  const uint8_t* img = p.masks + (size_t)b * p.slot_bytes;
  const size_t row_bytes = (size_t)p.wid * m;
  int32_t* cc = col_cnt + (size_t)b * p.wid * p.det_stride;  // [D][W]: detection-major, so the scan along x is contiguous
  uint32_t colbase[V];
  int n[V];
  int64_t out[V];
  W prevw = 0;
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int e = e0 + k;
    const int x = e / m, j = e - x * m;
    // predecessor of (y = 0, x): the last pixel of the previous column; of the very first pixel: 0
    const uint32_t pv = x > 0 ? img[(size_t)(p.hgt - 1) * row_bytes + e - m] : 0u;
    prevw |= (W)pv << (8 * k);
    colbase[k] = (uint32_t)x * (uint32_t)p.hgt;
    n[k] = 0;
    out[k] = EMIT ? pos_off[(size_t)b * p.det_stride + j] + cc[(size_t)j * p.wid + x] : 0;
  }
  const uint8_t* ptr = img + e0;
  constexpr int U = 8;
  W acc = 0;
  int since = 0;
  auto flush_acc = [&]() {
#pragma unroll
    for (int k = 0; k < V; ++k) n[k] += (int)((acc >> (8 * k)) & 0xffu);
    acc = 0;
    since = 0;
  };
  auto emit = [&](W f, int y) {
#pragma unroll
    for (int k = 0; k < V; ++k) {
      if ((f >> (8 * k)) & 0xffu) {
        if (out[k] + n[k] < cap) pos[out[k] + n[k]] = colbase[k] + (uint32_t)y;
        ++n[k];
      }
    }
  };
  int y = 0;
  for (; y + U <= p.hgt; y += U) {
    W w[U];
#pragma unroll
    for (int u = 0; u < U; ++u) w[u] = *reinterpret_cast<const W*>(ptr + (size_t)(y + u) * row_bytes);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const W f = w[u] ^ prevw;
      prevw = w[u];
      if (EMIT) {
        if (f) emit(f, y + u);
      } else {
        acc += f;
      }
    }
    if (!EMIT) {
      since += U;
      if (since > 255 - U) flush_acc();
    }
  }
  for (; y < p.hgt; ++y) {
    const W w = *reinterpret_cast<const W*>(ptr + (size_t)y * row_bytes);
    const W f = w ^ prevw;
    prevw = w;
    if (EMIT) {
      if (f) emit(f, y);
    } else {
      acc += f;
      if (++since >= 255) flush_acc();
    }
  }
  if (!EMIT) {
    flush_acc();
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const int e = e0 + k;
      const int x = e / m, j = e - x * m;
      cc[(size_t)j * p.wid + x] = n[k];
    }
  }
}

template <bool EMIT>
__global__ void __launch_bounds__(kRleThreads)
rle_walk_kernel(RleParams p, int32_t* __restrict__ col_cnt, const int64_t* __restrict__ pos_off, uint32_t* __restrict__ pos,
                int64_t cap) {
  const int b = blockIdx.y;
  const int m = rle_m(p, b);
  const int n_e = p.wid * m;
  const int t = blockIdx.x * kRleThreads + threadIdx.x;
  const bool al = ((p.slot_bytes | reinterpret_cast<uintptr_t>(p.masks)) & 7) == 0;
  if (al && (n_e & 7) == 0) {
    if (t * 8 < n_e) rle_walk<EMIT, 8>(p, b, m, t * 8, col_cnt, pos_off, pos, cap);
  } else if (al && (n_e & 3) == 0) {
    if (t * 4 < n_e) rle_walk<EMIT, 4>(p, b, m, t * 4, col_cnt, pos_off, pos, cap);
  } else if (al && (n_e & 1) == 0) {
    if (t * 2 < n_e) rle_walk<EMIT, 2>(p, b, m, t * 2, col_cnt, pos_off, pos, cap);
  } else {
    if (t < n_e) rle_walk<EMIT, 1>(p, b, m, t, col_cnt, pos_off, pos, cap);
  }
}

// B: one WARP per (image, detection): exclusive prefix of its column counts along x (contiguous in the [D][W] layout),
// total boundaries.
__global__ void __launch_bounds__(256) rle_scan_kernel(RleParams p, int32_t* __restrict__ col_cnt, int32_t* __restrict__ nb) {
  const int idx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (idx >= p.batch * p.det_stride) return;
  const int b = idx / p.det_stride, j = idx - b * p.det_stride;
  const int m = rle_m(p, b);
  if (j >= m) {
    if (lane == 0) nb[idx] = -1;  // no such detection
    return;
  }
  int32_t* cc = col_cnt + (size_t)b * p.wid * p.det_stride + (size_t)j * p.wid;
  int run = 0;
  for (int x0 = 0; x0 < p.wid; x0 += 32) {
    const int x = x0 + lane;
    const int c = x < p.wid ? cc[x] : 0;
    int incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (x < p.wid) cc[x] = run + incl - c;
    run += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) nb[idx] = run;
}

// B2: single CTA.  pos_off = exclusive prefix of nb over the valid detections, run_off likewise of (nb + 1) -- a
// detection with nb boundaries has nb + 1 runs.  run_count[idx] = nb + 1 (0 for missing detections).
__global__ void __launch_bounds__(1024) rle_offsets_kernel(const int32_t* __restrict__ nb, int n, int64_t* __restrict__ pos_off,
                                                           int64_t* __restrict__ run_off, int32_t* __restrict__ run_count,
                                                           int64_t* __restrict__ totals /* [2]: boundaries, runs */) {
  __shared__ int s_warp[32];
  __shared__ int64_t s_carry_b, s_carry_r;
  if (threadIdx.x == 0) { s_carry_b = 0; s_carry_r = 0; }
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < n ? nb[i] : -1;
    const int bcount = v >= 0 ? v : 0, rcount = v >= 0 ? v + 1 : 0;
    int tot_b, tot_r;
    const int eb = cta_exclusive_scan(bcount, s_warp, &tot_b);
    __syncthreads();
    const int er = cta_exclusive_scan(rcount, s_warp, &tot_r);
    if (i < n) {
      pos_off[i] = s_carry_b + eb;
      run_off[i] = s_carry_r + er;
      run_count[i] = rcount;
    }
    __syncthreads();
    if (threadIdx.x == 0) { s_carry_b += tot_b; s_carry_r += tot_r; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { totals[0] = s_carry_b; totals[1] = s_carry_r; }
}

// D: counts of detection idx: [pos0, pos1 - pos0, ..., HW - pos_last]  (or [HW] without boundaries)
__global__ void __launch_bounds__(kRleThreads)
rle_diff_kernel(const int32_t* __restrict__ nb, const int64_t* __restrict__ pos_off, const int64_t* __restrict__ run_off,
                const uint32_t* __restrict__ pos, uint32_t hw, uint32_t* __restrict__ counts, int64_t cap) {
  const int idx = blockIdx.x;
  const int n = nb[idx];
  if (n < 0) return;
  const int64_t po = pos_off[idx], ro = run_off[idx];
  for (int k = blockIdx.y * kRleThreads + threadIdx.x; k <= n; k += gridDim.y * kRleThreads) {
    if (ro + k >= cap) return;
    const uint32_t lo = k == 0 ? 0u : pos[po + k - 1];
    const uint32_t hi = k == n ? hw : pos[po + k];
    counts[ro + k] = hi - lo;
  }
}

int launch_rle(isb_handle* h, const uint8_t* masks, int64_t slot_bytes, const int32_t* counts_dev, int m_fixed, int det_stride,
               int batch, int hgt, int wid, uint32_t* counts, int64_t cap, int32_t* run_count, int64_t* run_off, int64_t* totals,
               cudaStream_t s) {
  // scratch: col_cnt int32 [B, D, W] | nb int32 [B*D] | pos_off int64 [B*D] | pos uint32 [cap]
  const size_t n_det = (size_t)batch * det_stride;
  const size_t b_col = ((size_t)batch * wid * det_stride * 4 + 255) & ~(size_t)255;
  const size_t b_nb = (n_det * 4 + 255) & ~(size_t)255;
  const size_t b_po = (n_det * 8 + 255) & ~(size_t)255;
  const size_t b_pos = ((size_t)cap * 4 + 255) & ~(size_t)255;
  const size_t need = b_col + b_nb + b_po + b_pos;
  if (need > h->rle_bytes) {
    ISB_CUDA(h, cudaStreamSynchronize(s));
    cudaFree(h->rle_scratch);
    h->rle_scratch = nullptr;
    h->rle_bytes = 0;
    ISB_CUDA(h, cudaMalloc(&h->rle_scratch, need));
    h->rle_bytes = need;
  }
  char* base = static_cast<char*>(h->rle_scratch);
  int32_t* col_cnt = reinterpret_cast<int32_t*>(base);
  int32_t* nb = reinterpret_cast<int32_t*>(base + b_col);
  int64_t* pos_off = reinterpret_cast<int64_t*>(base + b_col + b_nb);
  uint32_t* pos = reinterpret_cast<uint32_t*>(base + b_col + b_nb + b_po);
  RleParams p;
  p.masks = masks;
  p.slot_bytes = slot_bytes;
  p.counts = counts_dev;
  p.m_fixed = m_fixed;
  p.det_stride = det_stride;
  p.batch = batch;
  p.hgt = hgt;
  p.wid = wid;
  const dim3 walk_grid((unsigned)(((int64_t)wid * det_stride + kRleThreads - 1) / kRleThreads), (unsigned)batch);
  rle_walk_kernel<false><<<walk_grid, kRleThreads, 0, s>>>(p, col_cnt, nullptr, nullptr, 0);
  ISB_LAUNCH_CHECK(h, "rle_walk_kernel<count>");
  rle_scan_kernel<<<(unsigned)((n_det * 32 + 255) / 256), 256, 0, s>>>(p, col_cnt, nb);
  ISB_LAUNCH_CHECK(h, "rle_scan_kernel");
  rle_offsets_kernel<<<1, 1024, 0, s>>>(nb, (int)n_det, pos_off, run_off, run_count, totals);
  ISB_LAUNCH_CHECK(h, "rle_offsets_kernel");
  rle_walk_kernel<true><<<walk_grid, kRleThreads, 0, s>>>(p, col_cnt, pos_off, pos, cap);
  ISB_LAUNCH_CHECK(h, "rle_walk_kernel<emit>");
  rle_diff_kernel<<<dim3((unsigned)n_det, 8), kRleThreads, 0, s>>>(nb, pos_off, run_off, pos, (uint32_t)hgt * (uint32_t)wid, counts, cap);
  ISB_LAUNCH_CHECK(h, "rle_diff_kernel");
  return ISB_OK;
}

}  // namespace isb

extern "C" int isb_rle_encode(isb_handle* h, const uint8_t* masks, int64_t mask_slot_bytes, const int32_t* counts, int32_t m_fixed,
                              int32_t max_det, int32_t batch, int32_t hgt, int32_t wid, uint32_t* rle_counts, int64_t capacity,
                              int32_t* run_count, int64_t* run_offset, int64_t* totals, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (batch <= 0 || hgt <= 0 || wid <= 0 || max_det <= 0 || capacity <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "rle_encode: bad size");
  if (!masks || !rle_counts || !run_count || !run_offset || !totals) ISB_FAIL(h, ISB_ERR_BAD_ARG, "rle_encode: null pointer");
  if (!counts && (m_fixed < 0 || m_fixed > max_det)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "rle_encode: m_fixed must be in [0, max_det] without counts");
  if ((int64_t)hgt * wid >= (1ll << 32)) ISB_FAIL(h, ISB_ERR_LIMIT, "rle_encode: image too large");
  if ((int64_t)wid * max_det >= (1ll << 31) || (int64_t)batch * max_det >= (1ll << 31)) ISB_FAIL(h, ISB_ERR_LIMIT, "rle_encode: too many columns");
  if (batch > 1 && mask_slot_bytes < (int64_t)hgt * wid * max_det) ISB_FAIL(h, ISB_ERR_BAD_ARG, "rle_encode: mask_slot_bytes too small");
  ISB_SET_DEVICE(h);
  return isb::launch_rle(h, masks, mask_slot_bytes, counts, m_fixed, max_det, batch, hgt, wid, rle_counts, capacity, run_count,
                         run_offset, totals, (cudaStream_t)stream);
}

// ---- host: counts -> the ASCII string pycocotools puts into rle['counts'] (maskApi.c rleToString) -------------------
// Each count (from the 4th on: its difference to the count two places earlier) is written as a little-endian sequence of
// 5-bit groups, bit 0x20 = "more groups follow", offset by '0' (48); the value is sign-extended from the last group.
extern "C" int64_t isb_rle_to_string(const uint32_t* counts, int64_t n, char* out, int64_t out_cap) {
  if (n < 0 || (n > 0 && !counts) || !out) return -1;
  int64_t len = 0;
  for (int64_t i = 0; i < n; ++i) {
    long long x = (long long)counts[i];
    if (i > 2) x -= (long long)counts[i - 2];
    bool more = true;
    while (more) {
      char c = (char)(x & 0x1f);
      x >>= 5;
      more = (c & 0x10) ? x != -1 : x != 0;
      if (more) c |= 0x20;
      c += 48;
      if (len >= out_cap) return -2;
      out[len++] = c;
    }
  }
  return len;
}
