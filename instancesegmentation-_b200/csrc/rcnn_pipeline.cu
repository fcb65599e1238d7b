// Fused Mask R-CNN proposal / detection filters, one image per call, no host synchronisation:
//   isb_rpn_filter       = RegionProposalNetwork._get_top_n_idx + filter_proposals
//                          (rpn_function.py:434-458, :460-529)
//   isb_box_postprocess  = RoIHeads.postprocess_detections (roi_head.py:366-457)
// Both end in the batched NMS of nms.cu with a device-side box count.
#include <math.h>

#include "isb_common.cuh"
#include "nms_pair.cuh"

namespace isb {

struct LevelOffsets { int start[9]; int num_levels; };
struct LevelTopK { int k[8]; };  // min(pre_nms_top_n, level size)

// ---- RPN: exact per-level top-k by radix select (rpn_function.py:447-456) ------------------------------
// Every anchor has a unique 52-bit rank key  ~logit (32 bits, smaller = better) | anchor index (20 bits), i.e.
// score descending with ties by ascending index.  The k-th smallest key of each level is found digit by digit
// (11, 11, 11, 11, 8 bits): pass p histograms the digit of the anchors that match the prefix chosen so far; the
// next launch resolves that histogram (every CTA redundantly, 2048 bins) before it filters.  As soon as the chosen bin
// holds exactly the keys still wanted -- after the third digit at the latest when the logits are distinct, at once
// for a level with no more than k anchors -- the level is done and the remaining passes return immediately.  The objectness
// map (1 MB, L2-resident) is read six times; nothing is sorted until the exactly k survivors per level are
// known, so the full sort of 2^19 keys this replaces shrinks to one of 8192.
constexpr int kSelBins = 2048;
constexpr int kSelPasses = 5;
constexpr int kSelThreads = 256;
__device__ __constant__ int kSelShift[kSelPasses] = {41, 30, 19, 8, 0};
__device__ __constant__ int kSelBits[kSelPasses] = {11, 11, 11, 11, 8};

struct SelState { unsigned long long prefix; int k_rem; int done; };  // done: the chosen bin is taken whole, later digits are moot

__device__ __forceinline__ uint64_t rpn_rank_key(float logit, uint32_t index) {
  return ((uint64_t)float_desc_bits(logit) << kRowBits) | index;
}

// Resolve the histogram of pass `p` (p >= 0) given the state before it: pick the digit whose cumulative count
// reaches k_rem.  Called by all threads of a 256-thread CTA; the result is returned to every thread.
__device__ SelState sel_resolve(SelState st, const int* __restrict__ hist, int p, int* s_scan /* [32] */, int* s_res /* [3] */) {
  const int tid = threadIdx.x;
  if (st.done) {  // every key under the prefix is selected: the k-th key is the largest one under it
    SelState out;
    out.prefix = (st.prefix << kSelBits[p]) | (unsigned long long)((1 << kSelBits[p]) - 1);
    out.k_rem = st.k_rem;
    out.done = 1;
    return out;
  }
  int v[kSelBins / kSelThreads];
  int sum = 0;
#pragma unroll
  for (int i = 0; i < kSelBins / kSelThreads; ++i) { v[i] = hist[tid * (kSelBins / kSelThreads) + i]; sum += v[i]; }
  int total;
  int excl = cta_exclusive_scan(sum, s_scan, &total);
  if (excl < st.k_rem && st.k_rem <= excl + sum) {  // exactly one thread
    int b = 0, vb = 0;
#pragma unroll
    for (int i = 0; i < kSelBins / kSelThreads; ++i) {
      if (excl + v[i] >= st.k_rem) { b = i; vb = v[i]; break; }
      excl += v[i];
    }
    s_res[0] = tid * (kSelBins / kSelThreads) + b;
    s_res[1] = st.k_rem - excl;
    s_res[2] = vb == st.k_rem - excl;  // the bin holds exactly the keys still wanted (always true once keys are unique)
  }
  __syncthreads();
  SelState out;
  out.prefix = (st.prefix << kSelBits[p]) | (unsigned long long)s_res[0];
  out.k_rem = s_res[1];
  out.done = s_res[2];
  __syncthreads();
  return out;
}

// pass < kSelPasses: histogram the pass's digit.  pass == kSelPasses: emit the anchors whose key is <= the k-th key.
// grid (chunks, levels); hist [passes][levels][kSelBins]; state [passes + 1][levels] (state[0] = initial).
__global__ void __launch_bounds__(kSelThreads) rpn_select_pass_kernel(const float* __restrict__ objectness, LevelOffsets lv,
                                                                      LevelTopK topk, int pass, int* __restrict__ hist,
                                                                      SelState* __restrict__ state, int num_levels,
                                                                      int pre_top_n, uint64_t* __restrict__ cand_keys,
                                                                      int32_t* __restrict__ cand_count) {
  __shared__ int s_hist[kSelBins];
  __shared__ int s_scan[32];
  __shared__ int s_res[3];
  const int l = blockIdx.y, tid = threadIdx.x;
  const int begin = lv.start[l], end = lv.start[l + 1];
  if (begin >= end) return;
  // state[p] = prefix / remaining rank after the histogram of pass p - 1 has been resolved (written by launch p)
  SelState st;
  st.prefix = 0; st.k_rem = topk.k[l]; st.done = 0;
  if (pass > 0) {
    if (pass > 1) st = state[(size_t)(pass - 1) * num_levels + l];
    st = sel_resolve(st, hist + ((size_t)(pass - 1) * num_levels + l) * kSelBins, pass - 1, s_scan, s_res);
    if (blockIdx.x == 0 && tid == 0) state[(size_t)pass * num_levels + l] = st;
  }
  if (pass < kSelPasses) {
    if (st.done) return;  // nothing left to decide for this level: no histogram (the next launch fills the prefix with ones)
    for (int i = tid; i < kSelBins; i += kSelThreads) s_hist[i] = 0;
    __syncthreads();
    const int shift = kSelShift[pass], mask = (1 << kSelBits[pass]) - 1;
    for (int i = begin + blockIdx.x * kSelThreads + tid; i < end; i += gridDim.x * kSelThreads) {
      const uint64_t key = rpn_rank_key(objectness[i], (uint32_t)i);
      if (pass == 0 || (key >> (shift + kSelBits[pass])) == st.prefix) atomicAdd(&s_hist[(int)(key >> shift) & mask], 1);
    }
    __syncthreads();
    int* gh = hist + ((size_t)pass * num_levels + l) * kSelBins;
    for (int i = tid; i < kSelBins; i += kSelThreads)
      if (s_hist[i]) atomicAdd(&gh[i], s_hist[i]);
  } else {
    const uint64_t kth = st.prefix;  // all 52 bits resolved
    uint64_t* dst = cand_keys + (size_t)l * pre_top_n;
    for (int i = begin + blockIdx.x * kSelThreads + tid; i < end; i += gridDim.x * kSelThreads) {
      const float logit = objectness[i];
      if (rpn_rank_key(logit, (uint32_t)i) <= kth) {
        const int pos = atomicAdd(&cand_count[l], 1);
        dst[pos] = make_cand_key((uint32_t)l, logit, (uint32_t)i);  // exactly min(pre_top_n, level size) of these
      }
    }
  }
}

// candidate keys of all levels (unordered within a level) -> one padded array for the sort
__global__ void rpn_pack_keys_kernel(const uint64_t* __restrict__ cand_keys, LevelOffsets lv, int pre_top_n, int n_pow2,
                                     uint64_t* __restrict__ keys) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pow2; i += gridDim.x * blockDim.x) {
    uint64_t k = kPadKey;
    int rem = i;
    for (int l = 0; l < lv.num_levels; ++l) {
      const int kl = min(pre_top_n, lv.start[l + 1] - lv.start[l]);
      if (rem < kl) { k = cand_keys[(size_t)l * pre_top_n + rem]; break; }
      rem -= kl;
    }
    keys[i] = k;
  }
}

__device__ __forceinline__ int lower_bound_u64(const uint64_t* __restrict__ keys, int n, uint64_t v) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (keys[mid] < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ float4 clip_box(float4 b, float wf, float hf) {
  // boxes.py:108-138: x to [0, W], y to [0, H]
  b.x = fminf(fmaxf(b.x, 0.0f), wf);
  b.z = fminf(fmaxf(b.z, 0.0f), wf);
  b.y = fminf(fmaxf(b.y, 0.0f), hf);
  b.w = fminf(fmaxf(b.w, 0.0f), hf);
  return b;
}

// Single CTA: the levels' top-k in order (the sorted keys are level-major), clip, drop small / low-score boxes, compact
// in order.  A thread owns a run of consecutive candidates (at most kSelPerThread = ceil(total / 1024)): all its key and
// proposal loads are issued before anything depends on them, and ONE exclusive scan of the per-thread counts places
// every survivor -- two dependent global round trips and one scan instead of one of each per level.
constexpr int kSelPerThread = 8;  // 8 * 1024 candidates: 8 levels of pre_nms_top_n 1000; more falls back to a loop of passes

__global__ void __launch_bounds__(1024) rpn_select_kernel(const uint64_t* __restrict__ keys, LevelTopK topk,
                                                          const float4* __restrict__ proposals, int num_levels,
                                                          int pre_top_n, float img_w, float img_h, float min_size,
                                                          float score_thresh, float4* __restrict__ cbox,
                                                          float* __restrict__ cscore, int32_t* __restrict__ clvl,
                                                          int32_t* __restrict__ n_out) {
  __shared__ int s_scan[32];
  const int tid = threadIdx.x;
  int lvl_end[8];
  int total = 0;
#pragma unroll
  for (int l = 0; l < 8; ++l) {
    if (l < num_levels) total += topk.k[l];
    lvl_end[l] = total;  // candidates [lvl_end[l-1], lvl_end[l]) belong to level l
  }
  int base = 0;
  for (int p0 = 0; p0 < total; p0 += kSelPerThread * 1024) {  // one pass unless there are more than 8192 candidates
    const int n = min(total - p0, kSelPerThread * 1024);
    const int per = (n + 1023) / 1024;
    const int i0 = p0 + tid * per, i1 = min(i0 + per, p0 + n);
    uint64_t key[kSelPerThread];
    float4 bx[kSelPerThread];
#pragma unroll
    for (int q = 0; q < kSelPerThread; ++q)
      if (i0 + q < i1) key[q] = keys[i0 + q];
#pragma unroll
    for (int q = 0; q < kSelPerThread; ++q)
      if (i0 + q < i1) bx[q] = proposals[cand_row(key[q])];
    float sc[kSelPerThread];
    unsigned okmask = 0;
    int cnt = 0;
#pragma unroll
    for (int q = 0; q < kSelPerThread; ++q)
      if (i0 + q < i1) {
        bx[q] = clip_box(bx[q], img_w, img_h);                                                  // :509
        sc[q] = sigmoidf_ref(float_from_desc_bits(cand_score_bits(key[q])));                    // :502
        if ((bx[q].z - bx[q].x >= min_size) && (bx[q].w - bx[q].y >= min_size) && (sc[q] >= score_thresh)) {  // :512, :517
          okmask |= 1u << q;
          ++cnt;
        }
      }
    int pass_total;
    int slot = base + cta_exclusive_scan(cnt, s_scan, &pass_total);
#pragma unroll
    for (int q = 0; q < kSelPerThread; ++q)
      if ((okmask >> q) & 1u) {
        int l = 0;
        while (l < 7 && i0 + q >= lvl_end[l]) ++l;
        cbox[slot] = bx[q];
        cscore[slot] = sc[q];
        clvl[slot] = l;
        ++slot;
      }
    base += pass_total;
  }
  if (tid == 0) *n_out = base;
}

__global__ void rpn_gather_kernel(const int64_t* __restrict__ keep, const int32_t* __restrict__ count,
                                  const float4* __restrict__ cbox, const float* __restrict__ cscore, int max_out,
                                  float4* __restrict__ out_boxes, float* __restrict__ out_scores,
                                  int32_t* __restrict__ out_count) {
  const int n = min(*count, max_out);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int64_t k = keep[i];
    out_boxes[i] = cbox[k];
    out_scores[i] = cscore[k];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *out_count = n;
}

// ---- box head: one warp per proposal row ----------------------------------------------------
__device__ __forceinline__ float4 rcnn_decode_one(float4 rel, float4 b, float wx, float wy, float ww, float wh, float clip) {
  const float width = b.z - b.x, height = b.w - b.y;
  const float ctr_x = b.x + 0.5f * width, ctr_y = b.y + 0.5f * height;
  const float dx = rel.x / wx, dy = rel.y / wy;
  const float dw = fminf(rel.z / ww, clip), dh = fminf(rel.w / wh, clip);
  const float pcx = dx * width + ctr_x, pcy = dy * height + ctr_y;
  const float pw = expf(dw) * width, ph = expf(dh) * height;
  return make_float4(pcx - 0.5f * pw, pcy - 0.5f * ph, pcx + 0.5f * pw, pcy + 0.5f * ph);
}

struct BoxPostParams {
  const float* logits;   // [R, C] or nullptr
  const float* probs;    // [R, C] softmaxed, or nullptr (then logits are softmaxed here, roi_head.py:401)
  const float4* reg;     // [R, C] float4
  const float4* proposals;  // [R]
  int R, C;
  float wx, wy, ww, wh, clip, img_w, img_h, score_thresh;
  const int32_t* R_dev;  // optional: number of valid rows (<= R) in device memory; rows past it are ignored
};

__device__ __forceinline__ int boxpost_rows(const BoxPostParams& p) {
  if (!p.R_dev) return p.R;
  const int r = *p.R_dev;
  return r < 0 ? 0 : (r < p.R ? r : p.R);
}

// pass 0: count candidates per row (row_count); pass 1: write them at row_offset in class order.
template <int PASS>
__global__ void __launch_bounds__(256) boxpost_rows_kernel(BoxPostParams p, int32_t* __restrict__ row_count,
                                                           const int32_t* __restrict__ row_offset, int n_cap,
                                                           float4* __restrict__ cbox, float* __restrict__ cscore,
                                                           int32_t* __restrict__ clabel) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= p.R) return;
  if (r >= boxpost_rows(p)) {  // padded rows of a sync-free pipeline: no candidates
    if (!PASS && lane == 0) row_count[r] = 0;
    return;
  }
  float mx = -INFINITY, sum = 0.0f;
  if (!p.probs) {
    for (int c = lane; c < p.C; c += 32) mx = fmaxf(mx, p.logits[(size_t)r * p.C + c]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    for (int c = lane; c < p.C; c += 32) sum += expf(p.logits[(size_t)r * p.C + c] - mx);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  }
  const float4 prop = p.proposals[r];
  int written = 0;
  const int base = PASS ? row_offset[r] : 0;
  for (int c0 = 1; c0 < p.C; c0 += 32) {   // background class 0 dropped (roi_head.py:423-425)
    const int c = c0 + lane;
    bool ok = false;
    float sc = 0.f;
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < p.C) {
      sc = p.probs ? p.probs[(size_t)r * p.C + c] : expf(p.logits[(size_t)r * p.C + c] - mx) / sum;
      if (sc > p.score_thresh) {                                                         // :437
        b = clip_box(rcnn_decode_one(p.reg[(size_t)r * p.C + c], prop, p.wx, p.wy, p.ww, p.wh, p.clip), p.img_w, p.img_h);
        ok = (b.z - b.x >= 1.0f) && (b.w - b.y >= 1.0f);                                 // :442 min_size = 1
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, ok);
    if (PASS && ok) {
      const int slot = base + written + __popc(bal & ((1u << lane) - 1u));
      if (slot < n_cap) {
        cbox[slot] = b;
        cscore[slot] = sc;
        clabel[slot] = c;
      }
    }
    written += __popc(bal);
  }
  if (!PASS && lane == 0) row_count[r] = written;
}

// single CTA exclusive scan of the row counts; total (clamped to n_cap) -> n_out, overflow flag -> n_out[1]
__global__ void __launch_bounds__(1024) boxpost_scan_kernel(const int32_t* __restrict__ row_count, int R, int n_cap,
                                                            int32_t* __restrict__ row_offset, int32_t* __restrict__ n_out) {
  __shared__ int s_w[32];
  __shared__ int s_run;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_run = 0;
  __syncthreads();
  for (int r0 = 0; r0 < R; r0 += 1024) {
    const int r = r0 + tid;
    const int v = r < R ? row_count[r] : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    int woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 32; ++w) {
      const int c = s_w[w];
      if (w < warp) woff += c;
      tot += c;
    }
    if (r < R) row_offset[r] = s_run + woff + incl - v;
    __syncthreads();
    if (tid == 0) s_run += tot;
    __syncthreads();
  }
  if (tid == 0) {
    n_out[0] = min(s_run, n_cap);
    n_out[1] = s_run > n_cap ? 1 : 0;
  }
}

__global__ void boxpost_gather_kernel(const int64_t* __restrict__ keep, const int32_t* __restrict__ count,
                                      const int32_t* __restrict__ overflow, const float4* __restrict__ cbox,
                                      const float* __restrict__ cscore, const int32_t* __restrict__ clabel, int max_out,
                                      float4* __restrict__ out_boxes, float* __restrict__ out_scores,
                                      int64_t* __restrict__ out_labels, int32_t* __restrict__ out_count) {
  const int n = min(*count, max_out);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int64_t k = keep[i];
    out_boxes[i] = cbox[k];
    out_scores[i] = cscore[k];
    out_labels[i] = (int64_t)clabel[k];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *out_count = *overflow ? -1 : n;
}

static inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace isb

#include "rcnn_fused.cuh"

using namespace isb;

// ---- host side of the one-launch kernels (rcnn_fused.cuh) --------------------------------------------------------------
// Grid of the cooperative launch: every CTA must be resident at once, one 1024-thread CTA per SM.
template <class K>
static int fused_grid(isb_handle* h, int which, K kernel) {
  if (h->fused_grid[which] == 0) {
    const bool env_chain = getenv("ISB_RCNN_CHAIN") != nullptr;  // read once per handle: tests create a fresh one to get the chain
    int per_sm = 0, coop = 0;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFuSmemBytes);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kFuThreads, kFuSmemBytes);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device);
    if (e != cudaSuccess) cudaGetLastError();
    h->fused_grid[which] = (e == cudaSuccess && per_sm >= 1 && coop && !env_chain) ? h->num_sms : -1;
  }
  return h->fused_grid[which];
}

// carve the fused workspace out of the scratch: [zero block | extra_zero | arrays]; returns the bytes needed
static size_t fused_layout(char* base, int groups, size_t extra_zero, FusedWs* ws, char** extra, size_t* zero_bytes) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += up256(bytes); return p; };
  char* z = take(2 * 4 + (size_t)kFuMaxGroups * 4 + 2 * 4);
  char* x = take(extra_zero);
  const size_t zb = off;
  const size_t gc = (size_t)groups * kFuCap;
  char* lkey = take(gc * 8); char* lbox = take(gc * 16); char* sbox = take(gc * 16); char* sscore = take(gc * 4);
  char* spos = take(gc * 4); char* mask = take(gc * kFuWords * 8); char* kkey = take(gc * 8); char* kref = take(gc * 4);
  if (base) {
    ws->bar = (unsigned*)z;
    ws->gcount = (int32_t*)(z + 8);
    ws->gmax = (uint32_t*)(z + 8 + kFuMaxGroups * 4);
    ws->kept_total = (int32_t*)(z + 8 + kFuMaxGroups * 4 + 4);
    ws->lkey = (uint64_t*)lkey; ws->lbox = (float4*)lbox; ws->sbox = (float4*)sbox; ws->sscore = (float*)sscore;
    ws->spos = (uint32_t*)spos; ws->mask = (uint64_t*)mask; ws->kept_key = (uint64_t*)kkey; ws->kept_ref = (uint32_t*)kref;
    *extra = x;
    *zero_bytes = zb;
  }
  return off;
}

static float nms_threshold_f32(double thr) {  // largest float <= the double threshold: (double)iou > thr  <=>  iou > thr_f
  float t = (float)thr;
  if ((double)t > thr) t = nextafterf(t, -INFINITY);
  return t;
}

// ---- CUDA-graph replay of the two launch-bound chains ---------------------------------------------------------------
// isb_rpn_filter issues ~20 small kernels and isb_box_postprocess ~13; on a B200 their sum of kernel times is about two
// thirds of the wall time, the rest being launch gaps.  The first call with a given argument set runs eagerly (it also
// sizes the workspace); if the next call repeats it exactly -- same pointers, sizes and workspace, the steady state of a
// serving loop with preallocated tensors -- the sequence is captured on a private stream (nothing executes during
// capture) and from then on replayed with one cudaGraphLaunch into the caller's stream.  Any difference falls back to
// the eager path and restarts the cycle; ISB_NO_GRAPHS=1 disables it.
struct KeyBuilder {
  unsigned char buf[512];
  int len = 0;
  template <class T> void add(const T& v) {
    if (len + (int)sizeof(T) <= (int)sizeof(buf)) {
      memcpy(buf + len, &v, sizeof(T));
      len += (int)sizeof(T);
    } else {
      len = (int)sizeof(buf) + 1;  // too long: never matches
    }
  }
};

template <class F>
static int graph_run(isb_handle* h, int slot, const KeyBuilder& key, cudaStream_t s, F&& launch) {
  GraphSlot& g = h->graphs[slot];
  static const bool env_off = getenv("ISB_NO_GRAPHS") != nullptr;
  if (env_off || h->graphs_off || key.len > (int)sizeof(g.key)) return launch(s);
  if (g.exec && g.key_len == key.len && memcmp(g.key, key.buf, key.len) == 0) {
    ISB_CUDA(h, cudaGraphLaunch(g.exec, s));
    h->launches += g.nodes;
    return ISB_OK;
  }
  if (g.pending_len != key.len || memcmp(g.pending, key.buf, key.len) != 0) {  // first sighting: eager
    memcpy(g.pending, key.buf, key.len);
    g.pending_len = key.len;
    return launch(s);
  }
  // second consecutive call with these arguments: capture
  if (!h->capture_stream) ISB_CUDA(h, cudaStreamCreateWithFlags(&h->capture_stream, cudaStreamNonBlocking));
  if (g.exec) {
    cudaGraphExecDestroy(g.exec);
    g.exec = nullptr;
  }
  const int64_t launches0 = h->launches;
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamBeginCapture(h->capture_stream, cudaStreamCaptureModeThreadLocal);
  int rc = ISB_ERR_CUDA;
  if (e == cudaSuccess) {
    rc = launch(h->capture_stream);
    e = cudaStreamEndCapture(h->capture_stream, &graph);
  }
  const int nodes = (int)(h->launches - launches0);
  h->launches = launches0;
  if (e == cudaSuccess && rc == ISB_OK && graph) e = cudaGraphInstantiate(&g.exec, graph, 0);
  if (graph) cudaGraphDestroy(graph);
  if (e != cudaSuccess || rc != ISB_OK || !g.exec) {  // could not capture: stay eager from now on
    cudaGetLastError();
    g.exec = nullptr;
    h->graphs_off = 1;
    return launch(s);
  }
  memcpy(g.key, key.buf, key.len);
  g.key_len = key.len;
  g.nodes = nodes;
  ISB_CUDA(h, cudaGraphLaunch(g.exec, s));
  h->launches += g.nodes;
  return ISB_OK;
}

static void key_workspace(KeyBuilder& k, const isb_handle* h) {
  k.add(h->scratch); k.add(h->scratch_bytes); k.add(h->nms_cap); k.add(h->nms_keys); k.add(h->nms_keys2); k.add(h->nms_rank);
  k.add(h->nms_slabel); k.add(h->nms_sbox); k.add(h->nms_order); k.add(h->nms_mask); k.add(h->nms_scalars);
}

extern "C" int isb_rpn_filter(isb_handle* h, const float* proposals, const float* objectness,
                              const int32_t* anchors_per_level, int32_t num_levels, int32_t img_h, int32_t img_w,
                              int32_t pre_nms_top_n, int32_t post_nms_top_n, double nms_thresh, float score_thresh,
                              float min_size, float* out_boxes, float* out_scores, int32_t* out_count, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (!proposals || !objectness || !anchors_per_level || !out_boxes || !out_scores || !out_count)
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "rpn_filter: null pointer");
  if (num_levels < 1 || num_levels > 8 || pre_nms_top_n <= 0 || post_nms_top_n <= 0)
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "rpn_filter: 1..8 levels and positive top-n required");
  if ((reinterpret_cast<uintptr_t>(proposals) & 15) || (reinterpret_cast<uintptr_t>(out_boxes) & 15))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "rpn_filter: proposals / out_boxes must be 16-byte aligned");
  LevelOffsets lv;
  memset(&lv, 0, sizeof(lv));
  lv.num_levels = num_levels;
  int64_t total = 0, cand_max = 0;
  for (int l = 0; l < num_levels; ++l) {
    if (anchors_per_level[l] < 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "rpn_filter: negative level size");
    lv.start[l] = (int)total;
    total += anchors_per_level[l];
    cand_max += anchors_per_level[l] < pre_nms_top_n ? anchors_per_level[l] : pre_nms_top_n;
  }
  lv.start[num_levels] = (int)total;
  if (total > (1 << kRowBits)) ISB_FAIL(h, ISB_ERR_LIMIT, "rpn_filter: %lld anchors > 2^%d", (long long)total, kRowBits);
  ISB_SET_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  if (total == 0) {
    ISB_CUDA(h, cudaMemsetAsync(out_count, 0, sizeof(int32_t), s));
    return ISB_OK;
  }
  if (pre_nms_top_n <= kFuCap && nms_thresh >= 0.0) {
    const int grid = fused_grid(h, 0, rpn_filter_fused_kernel);
    if (grid >= num_levels) {
      const size_t b_hist = (size_t)kSelPasses * num_levels * kSelBins * 4;
      int rc = ensure_scratch(h, fused_layout(nullptr, num_levels, b_hist, nullptr, nullptr, nullptr));
      if (rc != ISB_OK) return rc;
      RpnFusedParams fp;
      memset(&fp, 0, sizeof(fp));
      char* hist = nullptr;
      size_t zero_bytes = 0;
      fused_layout(static_cast<char*>(h->scratch), num_levels, b_hist, &fp.ws, &hist, &zero_bytes);
      fp.objectness = objectness;
      fp.proposals = reinterpret_cast<const float4*>(proposals);
      fp.lv = lv;
      fp.num_levels = num_levels;
      fp.pre_top_n = pre_nms_top_n;
      fp.post_top_n = post_nms_top_n;
      fp.img_w = (float)img_w; fp.img_h = (float)img_h;
      fp.min_size = min_size; fp.score_thresh = score_thresh;
      fp.thr = nms_threshold_f32(nms_thresh);
      fp.hist = reinterpret_cast<int*>(hist);
      fp.out_boxes = reinterpret_cast<float4*>(out_boxes); fp.out_scores = out_scores; fp.out_count = out_count;
      // CTAs per level in the select passes, in proportion to the level's anchors (every non-empty level gets one)
      int given = 0, largest = 0;
      int nx[8];
      for (int l = 0; l < num_levels; ++l) {
        fp.topk.k[l] = anchors_per_level[l] < pre_nms_top_n ? anchors_per_level[l] : pre_nms_top_n;
        nx[l] = anchors_per_level[l] > 0 ? (int)((int64_t)(grid - num_levels) * anchors_per_level[l] / total) + 1 : 0;
        given += nx[l];
        if (anchors_per_level[l] > anchors_per_level[largest]) largest = l;
      }
      nx[largest] += grid - given;  // the rounding remainder goes to the largest level
      for (int l = 0; l < num_levels; ++l) fp.cta_begin[l + 1] = fp.cta_begin[l] + nx[l];
      ISB_CUDA(h, cudaMemsetAsync(h->scratch, 0, zero_bytes, s));
      void* args[] = {&fp};
      ISB_CUDA(h, cudaLaunchCooperativeKernel((void*)rpn_filter_fused_kernel, dim3(grid), dim3(kFuThreads), args, kFuSmemBytes, s));
      h->launches++;
      return ISB_OK;
    }
  }
  const int cmax = (int)cand_max, n_pow2 = next_pow2(cmax);
  const size_t b_keys = up256((size_t)n_pow2 * 8), b_box = up256((size_t)cmax * 16), b_sc = up256((size_t)cmax * 4),
               b_lvl = up256((size_t)cmax * 4), b_keep = up256((size_t)cmax * 8),
               b_cand = up256((size_t)num_levels * pre_nms_top_n * 8), b_sorted = up256((size_t)cmax * 8),
               b_rank = up256((size_t)cmax * 4),
               b_hist = up256((size_t)kSelPasses * num_levels * kSelBins * 4 + (size_t)num_levels * 4),
               b_state = up256((size_t)(kSelPasses + 1) * num_levels * sizeof(SelState));
  int rc = ensure_scratch(h, b_keys + b_box + b_sc + b_lvl + b_keep + b_cand + b_sorted + b_rank + b_hist + b_state + 512);
  if (rc != ISB_OK) return rc;
  if ((rc = ensure_nms_ws(h, cmax)) != ISB_OK) return rc;
  KeyBuilder key;
  key.add(proposals); key.add(objectness); key.add(num_levels); key.add(img_h); key.add(img_w); key.add(pre_nms_top_n);
  key.add(post_nms_top_n); key.add(nms_thresh); key.add(score_thresh); key.add(min_size); key.add(out_boxes); key.add(out_scores);
  key.add(out_count);
  for (int l = 0; l < num_levels; ++l) key.add(anchors_per_level[l]);
  key_workspace(key, h);
  return graph_run(h, 0, key, s, [&](cudaStream_t s) -> int {
    int rc;
    char* p = static_cast<char*>(h->scratch);
    uint64_t* keys = (uint64_t*)p;  p += b_keys;
    float4* cbox = (float4*)p;      p += b_box;
    float* cscore = (float*)p;      p += b_sc;
    int32_t* clvl = (int32_t*)p;    p += b_lvl;
    int64_t* keep = (int64_t*)p;    p += b_keep;
    uint64_t* cand = (uint64_t*)p;  p += b_cand;
    uint64_t* sorted = (uint64_t*)p; p += b_sorted;
    int32_t* rank = (int32_t*)p;    p += b_rank;
    int* hist = (int*)p;            // [passes][levels][bins] then cand_count [levels]
    int32_t* cand_count = hist + (size_t)kSelPasses * num_levels * kSelBins;
    p += b_hist;
    SelState* state = (SelState*)p; p += b_state;
    int32_t* n_dev = (int32_t*)p;   // [0] candidates, [2] kept count
    LevelTopK topk;
    memset(&topk, 0, sizeof(topk));
    int max_level = 0;
    for (int l = 0; l < num_levels; ++l) {
      topk.k[l] = anchors_per_level[l] < pre_nms_top_n ? anchors_per_level[l] : pre_nms_top_n;
      if (anchors_per_level[l] > max_level) max_level = anchors_per_level[l];
    }
    ISB_CUDA(h, cudaMemsetAsync(hist, 0, b_hist, s));
    int chunks = (max_level + kSelThreads * 8 - 1) / (kSelThreads * 8);
    if (chunks > 4 * h->num_sms) chunks = 4 * h->num_sms;
    if (chunks < 1) chunks = 1;
    for (int pass = 0; pass <= kSelPasses; ++pass) {
      rpn_select_pass_kernel<<<dim3(chunks, num_levels), kSelThreads, 0, s>>>(objectness, lv, topk, pass, hist, state, num_levels,
                                                                              pre_nms_top_n, cand, cand_count);
      ISB_LAUNCH_CHECK(h, "rpn_select_pass_kernel");
    }
    rpn_pack_keys_kernel<<<(n_pow2 + 255) / 256, 256, 0, s>>>(cand, lv, pre_nms_top_n, n_pow2, keys);
    ISB_LAUNCH_CHECK(h, "rpn_pack_keys_kernel");
    if (cmax <= kRankSortMax) {
      if ((rc = launch_rank_sort_u64(h, keys, cmax, nullptr, rank, sorted, s)) != ISB_OK) return rc;
    } else {
      if ((rc = launch_sort_u64(h, keys, n_pow2, s)) != ISB_OK) return rc;
      sorted = keys;
    }
    rpn_select_kernel<<<1, 1024, 0, s>>>(sorted, topk, reinterpret_cast<const float4*>(proposals), num_levels, pre_nms_top_n,
                                         (float)img_w, (float)img_h, min_size, score_thresh, cbox, cscore, clvl, n_dev);
    ISB_LAUNCH_CHECK(h, "rpn_select_kernel");
    NmsCall c;
    memset(&c, 0, sizeof(c));
    c.boxes = cbox;
    c.scores = cscore;
    c.idxs32 = clvl;
    c.n_max = cmax;
    c.n_dev = n_dev;
    c.iou_threshold = nms_thresh;
    c.max_keep = post_nms_top_n;
    c.keep = keep;
    c.count = n_dev + 2;
    if ((rc = launch_nms(h, c, s)) != ISB_OK) return rc;
    rpn_gather_kernel<<<8, 256, 0, s>>>(keep, n_dev + 2, cbox, cscore, post_nms_top_n, reinterpret_cast<float4*>(out_boxes),
                                        out_scores, out_count);
    ISB_LAUNCH_CHECK(h, "rpn_gather_kernel");
    return ISB_OK;
  });
}

static int box_postprocess_impl(isb_handle* h, const float* class_logits, const float* class_probs,
                                const float* box_regression, const float* proposals, int32_t num_rois,
                                const int32_t* num_rois_dev, int32_t num_classes, int32_t img_h, int32_t img_w, float wx,
                                float wy, float ww, float wh, float clip, float score_thresh, double nms_thresh,
                                int32_t detections_per_img, float* out_boxes, float* out_scores, int64_t* out_labels,
                                int32_t* out_count, isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if ((!class_logits && !class_probs) || !box_regression || !proposals || !out_boxes || !out_scores || !out_labels ||
      !out_count)
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "box_postprocess: null pointer");
  if (num_rois < 0 || num_classes < 2 || detections_per_img <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "box_postprocess: bad size");
  if ((reinterpret_cast<uintptr_t>(box_regression) & 15) || (reinterpret_cast<uintptr_t>(proposals) & 15) ||
      (reinterpret_cast<uintptr_t>(out_boxes) & 15))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "box_postprocess: box_regression / proposals / out_boxes must be 16-byte aligned");
  ISB_SET_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  if (num_rois == 0) {
    ISB_CUDA(h, cudaMemsetAsync(out_count, 0, sizeof(int32_t), s));
    return ISB_OK;
  }
  // probabilities of one row sum to 1, so fewer than 1/thresh classes can exceed the threshold
  int per_row = num_classes - 1;
  if (score_thresh > 0.0f) {
    const int lim = (int)floorf(1.0f / score_thresh);
    if (lim < per_row) per_row = lim > 1 ? lim : 1;
  }
  if (num_rois <= kFuCap && num_classes - 1 <= kFuMaxGroups && (int64_t)num_rois * num_classes < (1 << 20) && nms_thresh >= 0.0) {
    const int grid = fused_grid(h, 1, boxpost_fused_kernel);
    if (grid > 0) {
      const int groups = num_classes - 1;
      int rc = ensure_scratch(h, fused_layout(nullptr, groups, 0, nullptr, nullptr, nullptr));
      if (rc != ISB_OK) return rc;
      BoxFusedParams fp;
      memset(&fp, 0, sizeof(fp));
      char* unused = nullptr;
      size_t zero_bytes = 0;
      fused_layout(static_cast<char*>(h->scratch), groups, 0, &fp.ws, &unused, &zero_bytes);
      fp.bp.logits = class_probs ? nullptr : class_logits;
      fp.bp.probs = class_probs;
      fp.bp.reg = reinterpret_cast<const float4*>(box_regression);
      fp.bp.proposals = reinterpret_cast<const float4*>(proposals);
      fp.bp.R = num_rois;
      fp.bp.R_dev = num_rois_dev;
      fp.bp.C = num_classes;
      fp.bp.wx = wx; fp.bp.wy = wy; fp.bp.ww = ww; fp.bp.wh = wh; fp.bp.clip = clip;
      fp.bp.img_w = (float)img_w; fp.bp.img_h = (float)img_h;
      fp.bp.score_thresh = score_thresh;
      fp.thr = nms_threshold_f32(nms_thresh);
      fp.max_keep = detections_per_img;
      fp.out_boxes = reinterpret_cast<float4*>(out_boxes); fp.out_scores = out_scores; fp.out_labels = out_labels;
      fp.out_count = out_count;
      ISB_CUDA(h, cudaMemsetAsync(h->scratch, 0, zero_bytes, s));
      void* args[] = {&fp};
      ISB_CUDA(h, cudaLaunchCooperativeKernel((void*)boxpost_fused_kernel, dim3(grid), dim3(kFuThreads), args, kFuSmemBytes, s));
      h->launches++;
      return ISB_OK;
    }
  }
  const int64_t cap64 = (int64_t)num_rois * per_row;
  if (cap64 > (1 << 17)) ISB_FAIL(h, ISB_ERR_LIMIT, "box_postprocess: up to %lld candidates exceed the NMS limit", (long long)cap64);
  const int n_cap = (int)cap64;
  const size_t b_cnt = up256((size_t)num_rois * 4), b_box = up256((size_t)n_cap * 16), b_sc = up256((size_t)n_cap * 4),
               b_keep = up256((size_t)n_cap * 8);
  int rc = ensure_scratch(h, 2 * b_cnt + b_box + 2 * b_sc + b_keep + 512);
  if (rc != ISB_OK) return rc;
  if ((rc = ensure_nms_ws(h, n_cap)) != ISB_OK) return rc;
  KeyBuilder key;
  key.add(class_logits); key.add(class_probs); key.add(box_regression); key.add(proposals); key.add(num_rois); key.add(num_rois_dev);
  key.add(num_classes);
  key.add(img_h); key.add(img_w); key.add(wx); key.add(wy); key.add(ww); key.add(wh); key.add(clip); key.add(score_thresh);
  key.add(nms_thresh); key.add(detections_per_img); key.add(out_boxes); key.add(out_scores); key.add(out_labels); key.add(out_count);
  key_workspace(key, h);
  return graph_run(h, 1, key, s, [&](cudaStream_t s) -> int {
    int rc;
    char* p = static_cast<char*>(h->scratch);
    int32_t* row_count = (int32_t*)p;   p += b_cnt;
    int32_t* row_offset = (int32_t*)p;  p += b_cnt;
    float4* cbox = (float4*)p;          p += b_box;
    float* cscore = (float*)p;          p += b_sc;
    int32_t* clabel = (int32_t*)p;      p += b_sc;
    int64_t* keep = (int64_t*)p;        p += b_keep;
    int32_t* n_dev = (int32_t*)p;       // [0] candidates, [1] overflow, [2] kept count
    BoxPostParams bp;
    bp.logits = class_probs ? nullptr : class_logits;
    bp.probs = class_probs;
    bp.reg = reinterpret_cast<const float4*>(box_regression);
    bp.proposals = reinterpret_cast<const float4*>(proposals);
    bp.R = num_rois;
    bp.R_dev = num_rois_dev;
    bp.C = num_classes;
    bp.wx = wx; bp.wy = wy; bp.ww = ww; bp.wh = wh; bp.clip = clip;
    bp.img_w = (float)img_w; bp.img_h = (float)img_h;
    bp.score_thresh = score_thresh;
    const int row_blocks = (num_rois + 7) / 8;
    boxpost_rows_kernel<0><<<row_blocks, 256, 0, s>>>(bp, row_count, nullptr, n_cap, nullptr, nullptr, nullptr);
    ISB_LAUNCH_CHECK(h, "boxpost_rows_kernel<0>");
    boxpost_scan_kernel<<<1, 1024, 0, s>>>(row_count, num_rois, n_cap, row_offset, n_dev);
    ISB_LAUNCH_CHECK(h, "boxpost_scan_kernel");
    boxpost_rows_kernel<1><<<row_blocks, 256, 0, s>>>(bp, nullptr, row_offset, n_cap, cbox, cscore, clabel);
    ISB_LAUNCH_CHECK(h, "boxpost_rows_kernel<1>");
    NmsCall c;
    memset(&c, 0, sizeof(c));
    c.boxes = cbox;
    c.scores = cscore;
    c.idxs32 = clabel;
    c.n_max = n_cap;
    c.n_dev = n_dev;
    c.iou_threshold = nms_thresh;
    c.max_keep = detections_per_img;
    c.keep = keep;
    c.count = n_dev + 2;
    if ((rc = launch_nms(h, c, s)) != ISB_OK) return rc;
    boxpost_gather_kernel<<<4, 256, 0, s>>>(keep, n_dev + 2, n_dev + 1, cbox, cscore, clabel, detections_per_img,
                                            reinterpret_cast<float4*>(out_boxes), out_scores, out_labels, out_count);
    ISB_LAUNCH_CHECK(h, "boxpost_gather_kernel");
    return ISB_OK;
  });
}

extern "C" int isb_box_postprocess(isb_handle* h, const float* class_logits, const float* class_probs,
                                   const float* box_regression, const float* proposals, int32_t num_rois,
                                   int32_t num_classes, int32_t img_h, int32_t img_w, float wx, float wy, float ww,
                                   float wh, float clip, float score_thresh, double nms_thresh, int32_t detections_per_img,
                                   float* out_boxes, float* out_scores, int64_t* out_labels, int32_t* out_count,
                                   isb_stream stream) {
  return box_postprocess_impl(h, class_logits, class_probs, box_regression, proposals, num_rois, nullptr, num_classes, img_h, img_w,
                              wx, wy, ww, wh, clip, score_thresh, nms_thresh, detections_per_img, out_boxes, out_scores, out_labels,
                              out_count, stream);
}

extern "C" int isb_box_postprocess_rows(isb_handle* h, const float* class_logits, const float* class_probs,
                                        const float* box_regression, const float* proposals, int32_t max_rois,
                                        const int32_t* num_rois_dev, int32_t num_classes, int32_t img_h, int32_t img_w,
                                        float wx, float wy, float ww, float wh, float clip, float score_thresh,
                                        double nms_thresh, int32_t detections_per_img, float* out_boxes, float* out_scores,
                                        int64_t* out_labels, int32_t* out_count, isb_stream stream) {
  if (h && !num_rois_dev) ISB_FAIL(h, ISB_ERR_BAD_ARG, "box_postprocess_rows: num_rois_dev is null");
  return box_postprocess_impl(h, class_logits, class_probs, box_regression, proposals, max_rois, num_rois_dev, num_classes, img_h,
                              img_w, wx, wy, ww, wh, clip, score_thresh, nms_thresh, detections_per_img, out_boxes, out_scores,
                              out_labels, out_count, stream);
}
