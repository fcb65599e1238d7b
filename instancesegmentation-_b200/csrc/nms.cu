// Greedy NMS with torchvision semantics, and the batched (per-category) form.
//
// Replaces boxes.py:7-35 `nms` (-> torch.ops.torchvision.nms) and boxes.py:38-83 `batched_nms`.
// Semantics reproduced exactly (validated against the installed op, tests/golden/maskrcnn_stages):
//   * boxes are visited by descending score, ties by ascending index (the op sorts stably);
//   * a visited, unsuppressed box suppresses every later box with
//       inter / ((area_i + area_j) - inter) > iou_threshold,   all in float32,
//     the comparison being against the threshold as a DOUBLE (so the float threshold used here is the
//     largest float <= the double);
//   * batched_nms adds idx * (max_coordinate + 1) to every coordinate IN FLOAT32 before the IoU
//     (boxes.py:74-81), which costs mantissa bits -- reproduced, not "fixed".
//
// Pipeline (all on the caller's stream, no host sync):
//   keys (~score | index) -> bitonic sort -> gather boxes in score order (+offset) ->
//   64x64-tiled upper-triangular suppression bitmask (shared-memory column tile, one 64-bit word per
//   row and column block) -> single-CTA greedy scan over 64-box chunks (diagonal word resolved
//   serially from shared memory, kept rows OR-ed into the removed bitmap in parallel) ->
//   ordered compaction of the kept indices.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "isb_common.cuh"
#include "nms_pair.cuh"

namespace isb {

constexpr int kScanThreads = 1024;

__global__ void nms_max_coord_kernel(const float4* __restrict__ boxes, int n_max, const int32_t* __restrict__ n_dev,
                                     uint32_t* __restrict__ scalars) {
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  float m = -INFINITY;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 b = boxes[i];
    m = nan_max(nan_max(m, nan_max(b.x, b.y)), nan_max(b.z, b.w));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = nan_max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMin(&scalars[0], float_desc_bits(m));  // smaller desc-bits == larger float
}

// key = ~score | index; with `labels` (per-class NMS without coordinate offsets) the low word is
// label << 20 | index, so equal scores order by class first -- the order of the reference's per-class concatenation.
__global__ void nms_keys_kernel(const float* __restrict__ scores, const int32_t* __restrict__ labels, int n_max,
                                const int32_t* __restrict__ n_dev, int n_pow2, uint64_t* __restrict__ keys) {
  pdl_enter();
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pow2; i += gridDim.x * blockDim.x) {
    uint32_t low = (uint32_t)i;
    if (labels && i < n) low |= (uint32_t)labels[i] << kRowBits;
    keys[i] = i < n ? (((uint64_t)float_desc_bits(scores[i]) << 32) | low) : kPadKey;
  }
}

__global__ void nms_gather_kernel(const uint64_t* __restrict__ keys, const float4* __restrict__ boxes,
                                  const int64_t* __restrict__ idxs64, const int32_t* __restrict__ idxs32, int n_max,
                                  const int32_t* __restrict__ n_dev, const uint32_t* __restrict__ scalars,
                                  const int32_t* __restrict__ labels, float4* __restrict__ sbox,
                                  int32_t* __restrict__ order, int32_t* __restrict__ slabel) {
  pdl_enter();
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  const bool batched = idxs64 || idxs32;
  float step = 0.0f;
  if (batched) step = float_from_desc_bits(scalars[0]) + 1.0f;  // boxes.py:78  (max_coordinate + 1)
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
    const int i = labels ? (int)((uint32_t)keys[p] & (uint32_t)kRowMask) : (int)(uint32_t)keys[p];
    if (labels) slabel[p] = labels[i];
    else if (idxs32) slabel[p] = idxs32[i];  // level / class of the offset trick: lets the mask kernel skip cross-group pairs
    float4 b = boxes[i];
    if (batched) {
      const float off = (idxs64 ? (float)idxs64[i] : (float)idxs32[i]) * step;  // idxs.to(boxes) * (max + 1)
      b.x += off; b.y += off; b.z += off; b.w += off;                          // boxes + offsets[:, None]
    }
    sbox[p] = b;
    order[p] = i;
  }
}

// Suppression matrix, LOWER triangle (64x64 blocks with col_block <= row_block): bit j of mask[i][j / 64] is set iff
// i != j and IoU(i, j) > thr.  The scan only ever reads row i restricted to j < i ("who can suppress i"), so the blocks
// above the diagonal are neither computed nor launched (the words they would fill stay unwritten and unread).
// 1-D grid over the nb (nb + 1) / 2 blocks, 64 threads: thread t owns row rb * 64 + t against the 64 boxes of the column
// block staged in shared memory.  Disjoint pairs (the vast majority) skip the division: 0 / union is +-0 for every
// finite non-zero union, and the slow path of the IEEE division for a zero numerator is the most expensive
// thing in the loop.
__global__ void __launch_bounds__(64) nms_mask_kernel(const float4* __restrict__ sbox, int n_max,
                                                      const int32_t* __restrict__ n_dev, float thr, int words,
                                                      const int32_t* __restrict__ slabel, int offset_groups,
                                                      uint64_t* __restrict__ mask) {
  pdl_enter();
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  // linear block index -> (rb, cb), cb <= rb: rb = floor((sqrt(8 t + 1) - 1) / 2), fixed up for rounding
  const unsigned lin = blockIdx.x;
  int rb = (int)((sqrtf(8.0f * (float)lin + 1.0f) - 1.0f) * 0.5f);
  while ((unsigned)(rb + 1) * (unsigned)(rb + 2) / 2 <= lin) ++rb;
  while ((unsigned)rb * (unsigned)(rb + 1) / 2 > lin) --rb;
  const int cb = (int)(lin - (unsigned)rb * (unsigned)(rb + 1) / 2);
  if (rb * 64 >= n) return;
  __shared__ float4 cbox[64];
  __shared__ float carea[64];
  __shared__ int clabel[64];
  const int t = threadIdx.x;
  const int cj = cb * 64 + t;
  if (cj < n) {
    cbox[t] = sbox[cj];
    carea[t] = box_area(cbox[t]);
    clabel[t] = slabel ? slabel[cj] : 0;
  }
  __syncthreads();
  const int ri = rb * 64 + t;
  if (ri >= n) return;
  const float4 a = sbox[ri];
  const float aa = box_area(a);
  const int alabel = slabel ? slabel[ri] : 0;  // per-class mode: only boxes of the same class interact
  const int ncol = min(64, n - cb * 64);
  const bool zero_hits = 0.0f > thr;  // a zero IoU suppresses only for a negative threshold
  uint64_t bits = 0;
  for (int j = 0; j < ncol; ++j) {
    // boxes of different groups never overlap: per-class mode says so, and under the coordinate-offset trick
    // (boxes.py:70-82) their shifted copies are disjoint, so the IoU is exactly 0 -- no arithmetic needed.  (The
    // shortcut is only taken for the int32 group ids of the fused RPN pipeline, whose boxes are clipped to the image
    // first; with negative coordinates shifted groups CAN touch, and the int64 path of batched_nms computes every pair.)
    if (clabel[j] != alabel) {
      if (offset_groups && zero_hits) bits |= 1ull << j;
      continue;
    }
    // torchvision's CPU kernel semantics (NaN handling, division skipped for disjoint pairs): nms_pair.cuh
    const bool hit = nms_pair_hit(a, aa, cbox[j], carea[j], thr, zero_hits);
    if (hit) bits |= 1ull << j;
  }
  if (rb == cb) bits &= ~(1ull << t);  // a box does not suppress itself
  mask[(size_t)ri * words + cb] = bits;
}

// Single CTA: greedy NMS resolved 1024 boxes at a time by fixed-point iteration.
//   keep(i) = alive0(i) and no kept j < i of the chunk overlaps i,
//   alive0(i) = no kept box of an earlier chunk overlaps i.
// Starting from keep = alive0 the iteration alternates between over- and under-estimates of the greedy answer and
// settles after (length of the longest suppression chain) rounds, typically < 10; a round is 16 AND steps on
// register-resident words of the thread's own matrix row plus one ballot.  Thread t owns box chunk * 1024 + t.
// Replaces a 64-box-at-a-time scan that paid one global round trip and three barriers per 64 boxes.
constexpr int kScanChunkWords = kScanThreads / 64;  // 16

__global__ void __launch_bounds__(kScanThreads) nms_scan_kernel(const uint64_t* __restrict__ mask,
                                                                const int32_t* __restrict__ order, int n_max,
                                                                const int32_t* __restrict__ n_dev, int words, int max_keep,
                                                                int64_t* __restrict__ keep, int32_t* __restrict__ count) {
  pdl_enter();
  extern __shared__ uint64_t s_kept[];            // [words] kept bitmap over all boxes (score order)
  __shared__ uint32_t s_cur[kScanThreads / 32];   // this round's keep bits of the chunk, one ballot per warp
  __shared__ int s_warp_cnt[kScanThreads / 32];
  __shared__ int s_total;
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  const int nwords = (n + 63) >> 6;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int w = tid; w < nwords; w += kScanThreads) s_kept[w] = 0;
  if (tid == 0) s_total = 0;
  __syncthreads();
  const int nchunks = (n + kScanThreads - 1) / kScanThreads;
  for (int c = 0; c < nchunks; ++c) {
    const int i = c * kScanThreads + tid;
    const int w0 = c * kScanChunkWords;                 // first word of this chunk
    const int wn = min(kScanChunkWords, nwords - w0);    // words of this chunk
    const uint64_t* row = mask + (size_t)i * words;
    // alive0: not overlapped by any kept box of an earlier chunk
    bool alive = i < n;
    if (alive) {
      uint64_t acc = 0;
      for (int w = 0; w < w0; ++w) acc |= row[w] & s_kept[w];
      alive = acc == 0;
    }
    // own row inside the chunk, restricted to j < i
    uint64_t d[kScanChunkWords];
#pragma unroll
    for (int w = 0; w < kScanChunkWords; ++w) {
      uint64_t v = 0;
      if (alive && w < wn && w * 64 < tid) {
        v = row[w0 + w];
        if ((w + 1) * 64 > tid) v &= (1ull << (tid - w * 64)) - 1ull;
      }
      d[w] = v;
    }
    bool kp = alive;
    while (true) {
      const uint32_t bal = __ballot_sync(0xffffffffu, kp);
      if (lane == 0) s_cur[warp] = bal;
      __syncthreads();
      bool hit = false;
#pragma unroll
      for (int w = 0; w < kScanChunkWords; ++w) {
        const uint64_t kw = (uint64_t)s_cur[2 * w] | ((uint64_t)s_cur[2 * w + 1] << 32);
        hit |= (d[w] & kw) != 0;
      }
      const bool nk = alive && !hit;
      // the barrier also orders this round's reads of s_cur before the next round's writes
      const int changed = __syncthreads_or(nk != kp);
      kp = nk;
      if (!changed) break;
    }
    // count, cut at max_keep (in score order), publish
    const uint32_t bal = __ballot_sync(0xffffffffu, kp);
    if (lane == 0) s_warp_cnt[warp] = __popc(bal);
    __syncthreads();
    int before = s_total, chunk_total = 0;
    for (int w = 0; w < kScanThreads / 32; ++w) {
      const int cw = s_warp_cnt[w];
      if (w < warp) before += cw;
      chunk_total += cw;
    }
    const int my_pos = before + __popc(bal & ((1u << lane) - 1u));  // kept boxes before this one
    if (max_keep > 0 && my_pos >= max_keep) kp = false;
    const uint32_t bal2 = __ballot_sync(0xffffffffu, kp);
    if (lane == 0 && (warp >> 1) < wn) reinterpret_cast<uint32_t*>(s_kept + w0)[warp] = bal2;
    if (kp) keep[my_pos] = (int64_t)order[i];
    __syncthreads();
    if (tid == 0) s_total = (max_keep > 0 && s_total + chunk_total > max_keep) ? max_keep : s_total + chunk_total;
    __syncthreads();
    if (max_keep > 0 && s_total >= max_keep) break;
  }
  if (tid == 0) *count = s_total;
}

static void free_nms_ws(isb_handle* h) {
  cudaFree(h->nms_keys);
  cudaFree(h->nms_keys2);
  cudaFree(h->nms_rank);
  cudaFree(h->nms_slabel);
  cudaFree(h->nms_sbox);
  cudaFree(h->nms_order);
  cudaFree(h->nms_mask);
  h->nms_keys = nullptr; h->nms_keys2 = nullptr; h->nms_rank = nullptr; h->nms_slabel = nullptr; h->nms_sbox = nullptr; h->nms_order = nullptr; h->nms_mask = nullptr;
  h->nms_cap = 0;
}

int ensure_nms_ws(isb_handle* h, int64_t n) {
  if (!h->nms_scalars) ISB_CUDA(h, cudaMalloc((void**)&h->nms_scalars, 64));
  if (n <= h->nms_cap) return ISB_OK;
  if (n > (1 << 17)) ISB_FAIL(h, ISB_ERR_LIMIT, "nms over %lld boxes exceeds the 131072-box limit", (long long)n);
  ISB_CUDA(h, cudaDeviceSynchronize());
  free_nms_ws(h);
  int64_t cap = 8192;
  while (cap < n) cap <<= 1;
  const int64_t words = (cap + 63) / 64;
  ISB_CUDA(h, cudaMalloc((void**)&h->nms_keys, (size_t)cap * 8));
  ISB_CUDA(h, cudaMalloc((void**)&h->nms_keys2, (size_t)cap * 8));
  ISB_CUDA(h, cudaMalloc((void**)&h->nms_rank, (size_t)cap * 4));
  ISB_CUDA(h, cudaMalloc((void**)&h->nms_slabel, (size_t)cap * 4));
  ISB_CUDA(h, cudaMalloc((void**)&h->nms_sbox, (size_t)cap * 16));
  ISB_CUDA(h, cudaMalloc((void**)&h->nms_order, (size_t)cap * 4));
  ISB_CUDA(h, cudaMalloc((void**)&h->nms_mask, (size_t)cap * words * 8));
  h->nms_cap = cap;
  h->ws_bytes += (size_t)cap * (44 + words * 8);
  return ISB_OK;
}

int ensure_scratch(isb_handle* h, size_t bytes) {
  if (bytes <= h->scratch_bytes) return ISB_OK;
  ISB_CUDA(h, cudaDeviceSynchronize());
  cudaFree(h->scratch);
  h->scratch = nullptr;
  h->scratch_bytes = 0;
  ISB_CUDA(h, cudaMalloc(&h->scratch, bytes));
  h->scratch_bytes = bytes;
  return ISB_OK;
}

int configure_nms(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(nms_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  return ISB_OK;
}

int launch_nms(isb_handle* h, const NmsCall& c, cudaStream_t s) {
  int rc = ensure_nms_ws(h, c.n_max);
  if (rc != ISB_OK) return rc;
  if (c.n_max == 0) {
    ISB_CUDA(h, cudaMemsetAsync(c.count, 0, sizeof(int32_t), s));
    return ISB_OK;
  }
  const int n = c.n_max;
  const int n_pow2 = next_pow2(n);
  const int words = (n + 63) / 64;
  // largest float <= the double threshold: (double)ovr > thr  <=>  ovr > thr_f for float ovr
  float thr_f = (float)c.iou_threshold;
  if ((double)thr_f > c.iou_threshold) thr_f = nextafterf(thr_f, -INFINITY);
  const bool batched = c.idxs64 || c.idxs32;
  const int blocks = (n + 255) / 256 < 4 * h->num_sms ? (n + 255) / 256 : 4 * h->num_sms;
  // The memsets go first, so that the five to seven kernels behind them are adjacent on the stream and each can be the
  // programmatic dependent of the one before (pdl_enter / launch_chained): they are short, and the launch latency between
  // them was a large part of the call.
  const bool rank_sort = n <= kRankSortMax;
  if (rank_sort) ISB_CUDA(h, cudaMemsetAsync(h->nms_rank, 0, (size_t)n * sizeof(int32_t), s));
  cudaError_t le;
#define ISB_CHAINED(name, call)                                                                      \
  do {                                                                                               \
    le = (call);                                                                                     \
    if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "launch %s: %s", name, cudaGetErrorString(le)); \
    h->launches++;                                                                                   \
  } while (0)
  if (batched) {
    ISB_CUDA(h, cudaMemsetAsync(h->nms_scalars, 0xff, sizeof(uint32_t), s));
    nms_max_coord_kernel<<<blocks, 256, 0, s>>>(c.boxes, n, c.n_dev, h->nms_scalars);
    ISB_LAUNCH_CHECK(h, "nms_max_coord_kernel");
  }
  ISB_CHAINED("nms_keys_kernel", launch_chained(nms_keys_kernel, dim3((n_pow2 + 255) / 256), dim3(256), 0, s, batched, c.scores,
                                                c.labels32, n, c.n_dev, n_pow2, h->nms_keys));
  const uint64_t* sorted = h->nms_keys;
  if (rank_sort) {
    if ((rc = launch_rank_sort_u64(h, h->nms_keys, n, c.n_dev, h->nms_rank, h->nms_keys2, s, true)) != ISB_OK) return rc;
    sorted = h->nms_keys2;
  } else if ((rc = launch_sort_u64(h, h->nms_keys, n_pow2, s)) != ISB_OK) {
    return rc;
  }
  ISB_CHAINED("nms_gather_kernel", launch_chained(nms_gather_kernel, dim3(blocks), dim3(256), 0, s, rank_sort, sorted, c.boxes, c.idxs64,
                                                  c.idxs32, n, c.n_dev, (const uint32_t*)h->nms_scalars, c.labels32, h->nms_sbox,
                                                  h->nms_order, h->nms_slabel));
  ISB_CHAINED("nms_mask_kernel",
              launch_chained(nms_mask_kernel, dim3((unsigned)((int64_t)words * (words + 1) / 2)), dim3(64), 0, s, true,
                             (const float4*)h->nms_sbox, n, c.n_dev, thr_f, words,
                             (const int32_t*)((c.labels32 || c.idxs32) ? h->nms_slabel : nullptr), c.labels32 ? 0 : 1, h->nms_mask));
  const size_t smem = (size_t)words * 8;
  if (smem > 64 * 1024) ISB_FAIL(h, ISB_ERR_LIMIT, "nms scan shared memory");
  ISB_CHAINED("nms_scan_kernel", launch_chained(nms_scan_kernel, dim3(1), dim3(kScanThreads), smem, s, true, (const uint64_t*)h->nms_mask,
                                                (const int32_t*)h->nms_order, n, c.n_dev, words, c.max_keep, c.keep, c.count));
#undef ISB_CHAINED
  return ISB_OK;
}

}  // namespace isb

// ------------------------------------------------------------------------------------------------------------------
// isb_nms_static: the same greedy NMS for callers that may not allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
//
// Every scratch array of the pipeline above (sort keys, rank, boxes in score order, original indices, the suppression
// bit matrix) lives in ONE file-scope device array of a fixed, modest size; each GPU that loads libisb.so holds it
// until the process exits.  The array is shared by every call on that GPU and keeps the last call's contents: what
// must start at zero (the rank counters) is cleared on the caller's stream, everything else a call reads it has
// written first (keys incl. padding, order, boxes, the lower-triangle mask words).  Calls that overlap on the GPU
// (two streams, two threads) race on it.
namespace isb {
namespace {

constexpr int kStaticNmsMax = ISB_NMS_STATIC_MAX;              // boxes per call
constexpr int kStaticWords = kStaticNmsMax / 64;               // mask words per row at the maximum
constexpr size_t kStOffKeys = 0;                                                     // u64 [max]  (max is a power of two)
constexpr size_t kStOffKeys2 = kStOffKeys + (size_t)kStaticNmsMax * 8;               // u64 [max]
constexpr size_t kStOffBox = kStOffKeys2 + (size_t)kStaticNmsMax * 8;                // float4 [max]
constexpr size_t kStOffRank = kStOffBox + (size_t)kStaticNmsMax * 16;                // i32 [max]
constexpr size_t kStOffOrder = kStOffRank + (size_t)kStaticNmsMax * 4;               // i32 [max]
constexpr size_t kStOffMask = kStOffOrder + (size_t)kStaticNmsMax * 4;               // u64 [max][max / 64]
constexpr size_t kStaticNmsBytes = kStOffMask + (size_t)kStaticNmsMax * kStaticWords * 8;  // 2 260 992 bytes at 4096
static_assert((kStaticNmsMax & (kStaticNmsMax - 1)) == 0 && kStaticNmsMax <= kRankSortMax, "power of two within the rank sort");
static_assert(kStOffMask % 256 == 0, "every sub-array stays 256-byte aligned");

// In HOST code this name is only a stand-in for the symbol: its device address comes from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_nms_static_scratch[kStaticNmsBytes];

constexpr int kStaticMaxDevices = 64;
struct StaticNmsDevice {
  std::once_flag once;
  cudaError_t status = cudaErrorUnknown;
  unsigned char* base = nullptr;  // device address of g_nms_static_scratch on this GPU
  int num_sms = 0;
};
StaticNmsDevice g_static_dev[kStaticMaxDevices];

// once per device, with that device current
void static_nms_setup(StaticNmsDevice* d, int device) {
  void* p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, g_nms_static_scratch);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&d->num_sms, cudaDevAttrMultiProcessorCount, device);
  d->base = static_cast<unsigned char*>(p);
  d->status = e;
}

int static_fail(char* err, int32_t err_len, int code, const char* msg) {
  if (err && err_len > 0) snprintf(err, (size_t)err_len, "%s", msg);
  return code;
}

}  // namespace
}  // namespace isb

extern "C" int isb_nms_static(const float* boxes, const float* scores, int32_t n, double iou_threshold, int64_t* keep,
                              int32_t* count, isb_stream stream, char* err, int32_t err_len) {
  using namespace isb;
  char msg[256];
  if (err && err_len > 0) err[0] = 0;
  if (n < 0 || !count) return static_fail(err, err_len, ISB_ERR_BAD_ARG, "nms_static: negative n or null count");
  if (n > kStaticNmsMax) {
    snprintf(msg, sizeof(msg), "nms_static: %d boxes exceed the fixed scratch of ISB_NMS_STATIC_MAX = %d boxes (use isb_nms)", n,
             kStaticNmsMax);
    return static_fail(err, err_len, ISB_ERR_LIMIT, msg);
  }
  cudaStream_t s = (cudaStream_t)stream;
  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess || device < 0 || device >= kStaticMaxDevices) {
    snprintf(msg, sizeof(msg), "nms_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), device);
    return static_fail(err, err_len, ISB_ERR_CUDA, msg);
  }
  StaticNmsDevice* d = &g_static_dev[device];
  std::call_once(d->once, static_nms_setup, d, device);
  if (d->status != cudaSuccess || !d->base) {
    snprintf(msg, sizeof(msg), "nms_static: cudaGetSymbolAddress(scratch) on device %d: %s", device, cudaGetErrorString(d->status));
    return static_fail(err, err_len, ISB_ERR_CUDA, msg);
  }
  if (n == 0) {  // boxes.py:66-67: empty input -> empty result
    e = cudaMemsetAsync(count, 0, sizeof(int32_t), s);
    if (e != cudaSuccess) return static_fail(err, err_len, ISB_ERR_CUDA, cudaGetErrorString(e));
    return ISB_OK;
  }
  if (!boxes || !scores || !keep || (reinterpret_cast<uintptr_t>(boxes) & 15))
    return static_fail(err, err_len, ISB_ERR_BAD_ARG, "nms_static: null or misaligned pointer");

  uint64_t* keys = reinterpret_cast<uint64_t*>(d->base + kStOffKeys);
  uint64_t* keys2 = reinterpret_cast<uint64_t*>(d->base + kStOffKeys2);
  float4* sbox = reinterpret_cast<float4*>(d->base + kStOffBox);
  int32_t* rank = reinterpret_cast<int32_t*>(d->base + kStOffRank);
  int32_t* order = reinterpret_cast<int32_t*>(d->base + kStOffOrder);
  uint64_t* mask = reinterpret_cast<uint64_t*>(d->base + kStOffMask);

  // launch geometry and error text of the shared launchers come from a handle: a plain stack object stands in (no workspace)
  isb_handle ctx;
  memset(&ctx, 0, sizeof(ctx));
  ctx.device = device;
  ctx.num_sms = d->num_sms;
  const int n_pow2 = next_pow2(n);
  const int words = (n + 63) / 64;
  float thr_f = (float)iou_threshold;  // largest float <= the double threshold, as in launch_nms
  if ((double)thr_f > iou_threshold) thr_f = nextafterf(thr_f, -INFINITY);
  const int blocks = (n + 255) / 256 < 4 * d->num_sms ? (n + 255) / 256 : 4 * d->num_sms;
  int rc = ISB_OK;
  const char* where = "";
  do {
    nms_keys_kernel<<<(n_pow2 + 255) / 256, 256, 0, s>>>(scores, nullptr, n, nullptr, n_pow2, keys);
    if ((e = cudaGetLastError()) != cudaSuccess) { where = "nms_keys_kernel"; break; }
    if ((rc = launch_rank_sort_u64(&ctx, keys, n, nullptr, rank, keys2, s)) != ISB_OK) break;  // clears rank[0..n) first
    nms_gather_kernel<<<blocks, 256, 0, s>>>(keys2, reinterpret_cast<const float4*>(boxes), nullptr, nullptr, n, nullptr, nullptr,
                                             nullptr, sbox, order, nullptr);
    if ((e = cudaGetLastError()) != cudaSuccess) { where = "nms_gather_kernel"; break; }
    nms_mask_kernel<<<(unsigned)((int64_t)words * (words + 1) / 2), 64, 0, s>>>(sbox, n, nullptr, thr_f, words, nullptr, 1, mask);
    if ((e = cudaGetLastError()) != cudaSuccess) { where = "nms_mask_kernel"; break; }
    nms_scan_kernel<<<1, kScanThreads, (size_t)words * 8, s>>>(mask, order, n, nullptr, words, 0, keep, count);
    if ((e = cudaGetLastError()) != cudaSuccess) { where = "nms_scan_kernel"; break; }
  } while (0);
  if (rc != ISB_OK) return static_fail(err, err_len, rc, ctx.err);
  if (e != cudaSuccess) {
    snprintf(msg, sizeof(msg), "nms_static: launch %s: %s", where, cudaGetErrorString(e));
    return static_fail(err, err_len, ISB_ERR_CUDA, msg);
  }
  return ISB_OK;
}
