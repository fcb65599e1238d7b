// One cooperative launch per call for the two Mask R-CNN filters (included by rcnn_pipeline.cu, same translation unit).
//
// isb_rpn_filter and isb_box_postprocess were chains of ~20 / ~13 small dependent kernels: ~137 / ~95 us of kernel time for
// 14 MB / 0.5 MB of algorithmic traffic, almost all of it launch boundaries, ramp-up and tails (profiles/r02_maskrcnn_launches.csv).
// Here the whole call is ONE persistent kernel of one 1024-thread CTA per SM; the dependent steps are separated by a grid
// barrier (~1.5 us) instead of a kernel boundary, and two algorithmic observations shorten the steps themselves:
//
//  * batched_nms (boxes.py:38-83) on boxes that were clipped to the image first is an independent greedy NMS per group
//    (FPN level / class): shifted copies of different groups cannot overlap, so for a threshold >= 0 they never suppress
//    each other.  A group holds at most 1024 candidates (pre_nms_top_n per level; one candidate per proposal row and
//    class), so ONE CTA sorts a group in shared memory (bitonic network) and ONE CTA runs its greedy scan as a single
//    1024-box chunk, all groups in parallel; the suppression bits of the groups' lower triangles are spread over every
//    warp of the grid.  The coordinate offsets idx * (max + 1) are still ADDED in float32 before the IoU, because they
//    change the low mantissa bits of the coordinates -- reproduced, not "fixed".
//  * the order the reference's flattened candidate list would have is carried in the sort keys (level, rank in level) /
//    (row * C + class), so nothing has to be compacted in order between the steps; the final result order
//    (score descending, ties by list position = torch's stable sort inside torchvision nms) is one all-pairs rank over the
//    kept boxes, a warp per kept box.
//
// The chain stays as the general path (more than 1024 candidates per group, more than 128 groups, a negative IoU
// threshold -- where disjoint boxes DO suppress each other) and ISB_RCNN_CHAIN=1 forces it for A/B runs.
#pragma once

namespace isb {

constexpr int kFuThreads = 1024;
constexpr int kFuWarps = kFuThreads / 32;
constexpr int kFuCap = 1024;             // candidates per group
constexpr int kFuWords = kFuCap / 64;    // suppression words per matrix row
constexpr int kFuMaxGroups = 128;
constexpr int kFuMergeTile = 4096;       // kept keys staged per tile of the final rank
constexpr int kFuSmemBytes = kFuWarps * 64 * 20;  // largest phase: per-warp column tiles (float4 box + float area) = 40 KB

// device pointers into the handle's scratch; the first `zero_bytes` bytes of the block are cleared by one memset per call
struct FusedWs {
  unsigned* bar;        // [2] grid barrier: arrivals, generation
  int32_t* gcount;      // [groups] P0: candidates appended per group; P1 on: valid boxes of the group in rank order
  uint32_t* gmax;       // [1] ~float_desc_bits(max coordinate of the valid boxes) (atomicMax, 0 = none)
  int32_t* kept_total;  // [1]
  uint64_t* lkey;       // [groups][cap] candidate keys as they were found (unordered)
  float4* lbox;         // [groups][cap] box head: candidate boxes beside the keys
  float4* sbox;         // [groups][cap] boxes in rank order, clipped, NOT shifted
  float* sscore;        // [groups][cap]
  uint32_t* spos;       // [groups][cap] position in the reference's flattened candidate list (tie-break of the result order)
  uint64_t* mask;       // [groups][cap][cap / 64] suppression bits, lower triangle
  uint64_t* kept_key;   // [groups * cap] ~score | pos of the kept boxes, unordered
  uint32_t* kept_ref;   // [groups * cap] group * cap + rank
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// All CTAs of a cooperative launch.  Generation counter: the last arrival clears the count and releases the next
// generation; the acquire load that sees it orders every CTA's earlier global writes before the readers' later loads.
// `gen` is the caller's count of barriers passed so far (the words are zeroed before the launch, so the generation a CTA
// waits to leave is known without reading it: one L2 round trip less per barrier).
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned nblocks, unsigned& gen) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(bar, 1u) == nblocks - 1) {
      bar[0] = 0;
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(bar + 1), "r"(gen + 1) : "memory");
    } else {
      while (ld_acquire_u32(bar + 1) == gen) {}
    }
  }
  ++gen;
  __syncthreads();
}

// ascending bitonic sort of s[0 .. len) (len a power of two <= 1024) by the whole CTA; ends with a barrier
__device__ __forceinline__ void cta_bitonic(uint64_t* s, int len) {
  const int t = threadIdx.x;
  for (int size = 2; size <= len; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      if (t < (len >> 1)) {
        const int lo = 2 * t - (t & (stride - 1));
        const uint64_t a = s[lo], b = s[lo + stride];
        if ((a > b) == ((lo & size) == 0)) { s[lo] = b; s[lo + stride] = a; }
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ int pow2_at_least(int n) {
  int p = 2;
  while (p < n) p <<= 1;
  return p;
}

// max over the valid threads' boxes into ws.gmax (NaN propagates, as torch's boxes.max()); called by full warps
__device__ __forceinline__ void fused_max_coord(bool ok, float4 b, uint32_t* gmax) {
  float m = ok ? nan_max(nan_max(b.x, b.y), nan_max(b.z, b.w)) : -INFINITY;
  const unsigned any = __ballot_sync(0xffffffffu, ok);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = nan_max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (any && (threadIdx.x & 31) == 0) atomicMax(gmax, ~float_desc_bits(m));
}

// ---- P2: suppression bits of every group's lower triangle, a warp per (32 rows x 64 columns) item -------------------
// id0: the reference's idxs value of group 0 (levels count from 0, foreground classes from 1): offset = idx * (max + 1).
__device__ void fused_mask_phase(const FusedWs& ws, int groups, int id0, float thr, unsigned char* smem, int* s_n /*[groups]*/,
                                 int* s_start /*[groups + 1]*/, int* s_scan /*[32]*/) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int items = 0;
  if (tid < groups) {
    const int n = __ldcg(ws.gcount + tid);
    s_n[tid] = n;
    items = ((n + 31) >> 5) * ((n + 63) >> 6);  // rectangular; items above the diagonal return at once
  }
  int total;
  const int excl = cta_exclusive_scan(items, s_scan, &total);
  if (tid < groups) s_start[tid] = excl;
  if (tid == 0) s_start[groups] = total;
  __syncthreads();
  const float step = float_from_desc_bits(~__ldcg(ws.gmax)) + 1.0f;  // boxes.py:78  (max_coordinate + 1)
  float4* cbox = reinterpret_cast<float4*>(smem) + warp * 64;
  float* carea = reinterpret_cast<float*>(smem + (size_t)kFuWarps * 64 * 16) + warp * 64;
  const bool zero_hits = 0.0f > thr;
  for (int item = blockIdx.x * kFuWarps + warp; item < total; item += gridDim.x * kFuWarps) {
    int g = 0;
    {  // group of the item: largest g with s_start[g] <= item
      int lo = 0, hi = groups;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_start[mid] <= item) lo = mid; else hi = mid;
      }
      g = lo;
    }
    const int n = s_n[g], nb = (n + 63) >> 6;
    const int local = item - s_start[g];
    const int r32 = local / nb, cb = local - r32 * nb;
    if (cb > (r32 >> 1)) continue;  // above the diagonal
    const float off = (float)(g + id0) * step;   // idxs.to(boxes) * (max + 1)
    const float4* gbox = ws.sbox + (size_t)g * kFuCap;
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int cj = cb * 64 + q * 32 + lane;
      if (cj < n) {
        float4 b = __ldcg(gbox + cj);
        b.x += off; b.y += off; b.z += off; b.w += off;   // boxes + offsets[:, None]
        cbox[q * 32 + lane] = b;
        carea[q * 32 + lane] = box_area(b);
      }
    }
    __syncwarp();
    const int ri = r32 * 32 + lane;
    if (ri < n) {
      float4 a = __ldcg(gbox + ri);
      a.x += off; a.y += off; a.z += off; a.w += off;
      const float aa = box_area(a);
      const int ncol = min(64, n - cb * 64);
      uint64_t bits = 0;
      for (int j = 0; j < ncol; ++j)
        if (nms_pair_hit(a, aa, cbox[j], carea[j], thr, zero_hits)) bits |= 1ull << j;
      if (cb == (ri >> 6)) bits &= ~(1ull << (ri & 63));  // a box does not suppress itself
      ws.mask[((size_t)g * kFuCap + ri) * kFuWords + cb] = bits;
    }
  }
}

// ---- P3: greedy scan of a group by one CTA (the single-chunk case of nms_scan_kernel), kept boxes appended to the kept list
__device__ void fused_scan_phase(const FusedWs& ws, int groups, int max_keep, uint32_t* s_cur /*[32]*/, int* s_cnt /*[32]*/,
                                 int* s_base) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int g = blockIdx.x; g < groups; g += gridDim.x) {
    const int n = __ldcg(ws.gcount + g);
    if (n <= 0) continue;  // CTA-uniform
    const uint64_t* row = ws.mask + ((size_t)g * kFuCap + tid) * kFuWords;
    const bool alive = tid < n;
    uint64_t d[kFuWords];  // own row restricted to j < tid
#pragma unroll
    for (int w = 0; w < kFuWords; ++w) {
      uint64_t v = 0;
      if (alive && w * 64 < tid) {
        v = __ldcg(row + w);
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
      for (int w = 0; w < kFuWords; ++w) {
        const uint64_t kw = (uint64_t)s_cur[2 * w] | ((uint64_t)s_cur[2 * w + 1] << 32);
        hit |= (d[w] & kw) != 0;
      }
      const bool nk = alive && !hit;
      const int changed = __syncthreads_or(nk != kp);  // also orders this round's reads of s_cur before the next writes
      kp = nk;
      if (!changed) break;
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, kp);
    if (lane == 0) s_cnt[warp] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < kFuWarps; ++w) {
      const int c = s_cnt[w];
      if (w < warp) before += c;
      total += c;
    }
    const int my_pos = before + __popc(bal & ((1u << lane) - 1u));
    const int emit = min(total, max_keep);  // kept boxes of a group are in result order already: later ones cannot make the cut
    if (tid == 0) *s_base = atomicAdd(ws.kept_total, emit);
    __syncthreads();
    if (kp && my_pos < emit) {
      const size_t at = (size_t)g * kFuCap + tid;
      const int slot = *s_base + my_pos;
      ws.kept_key[slot] = ((uint64_t)float_desc_bits(__ldcg(ws.sscore + at)) << 32) | __ldcg(ws.spos + at);
      ws.kept_ref[slot] = (uint32_t)at;
    }
    __syncthreads();
  }
}

// ---- P4: result order = score descending, ties by list position; a warp ranks one kept box against all of them --------
__device__ void fused_merge_phase(const FusedWs& ws, int max_keep, int id0, unsigned char* smem, float4* out_boxes, float* out_scores,
                                  int64_t* out_labels, int32_t* out_count) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kept = __ldcg(ws.kept_total);
  uint64_t* s_key = reinterpret_cast<uint64_t*>(smem);  // [kFuMergeTile]
  const int gw = blockIdx.x * kFuWarps + warp, nw = gridDim.x * kFuWarps;
  const int rounds = (kept + nw - 1) / nw;  // CTA-uniform trip count: the tile loads below need every thread
  for (int r = 0; r < rounds; ++r) {
    const int e = gw + r * nw;
    const uint64_t mine = e < kept ? __ldcg(ws.kept_key + e) : 0ull;
    int cnt = 0;
    for (int t0 = 0; t0 < kept; t0 += kFuMergeTile) {
      const int tn = min(kFuMergeTile, kept - t0);
      __syncthreads();
      for (int i = tid; i < tn; i += kFuThreads) s_key[i] = __ldcg(ws.kept_key + t0 + i);
      __syncthreads();
      if (e < kept)
        for (int i = lane; i < tn; i += 32) cnt += s_key[i] < mine ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (e < kept && cnt < max_keep && lane == 0) {
      const uint32_t at = __ldcg(ws.kept_ref + e);
      out_boxes[cnt] = __ldcg(ws.sbox + at);
      out_scores[cnt] = __ldcg(ws.sscore + at);
      if (out_labels) out_labels[cnt] = (int64_t)(at / kFuCap) + id0;
    }
  }
  if (blockIdx.x == 0 && tid == 0) *out_count = min(kept, max_keep);
}

// ======================================================================================================================
// isb_rpn_filter in one launch
// ======================================================================================================================
struct RpnFusedParams {
  const float* objectness;
  const float4* proposals;
  LevelOffsets lv;
  LevelTopK topk;
  int cta_begin[9];  // CTAs [cta_begin[l], cta_begin[l + 1]) read level l during the radix select
  int num_levels, pre_top_n, post_top_n;
  float img_w, img_h, min_size, score_thresh, thr;
  int* hist;         // [passes][levels][kSelBins], cleared by the per-call memset
  FusedWs ws;
  float4* out_boxes;
  float* out_scores;
  int32_t* out_count;
};

// sel_resolve for a 1024-thread CTA (two bins per thread), histogram read past L1 (other CTAs' atomics filled it)
__device__ SelState sel_resolve_1024(SelState st, const int* hist, int p, int* s_scan, int* s_res) {
  const int tid = threadIdx.x;
  if (st.done) {
    SelState out;
    out.prefix = (st.prefix << kSelBits[p]) | (unsigned long long)((1 << kSelBits[p]) - 1);
    out.k_rem = st.k_rem;
    out.done = 1;
    return out;
  }
  constexpr int kPer = kSelBins / kFuThreads;  // 2
  int v[kPer];
  int sum = 0;
#pragma unroll
  for (int i = 0; i < kPer; ++i) { v[i] = __ldcg(hist + tid * kPer + i); sum += v[i]; }
  int total;
  int excl = cta_exclusive_scan(sum, s_scan, &total);
  if (excl < st.k_rem && st.k_rem <= excl + sum) {  // exactly one thread
    int b = 0, vb = 0;
#pragma unroll
    for (int i = 0; i < kPer; ++i) {
      if (excl + v[i] >= st.k_rem) { b = i; vb = v[i]; break; }
      excl += v[i];
    }
    s_res[0] = tid * kPer + b;
    s_res[1] = st.k_rem - excl;
    s_res[2] = vb == st.k_rem - excl;
  }
  __syncthreads();
  SelState out;
  out.prefix = (st.prefix << kSelBits[p]) | (unsigned long long)s_res[0];
  out.k_rem = s_res[1];
  out.done = s_res[2];
  __syncthreads();
  return out;
}

__global__ void __launch_bounds__(kFuThreads, 1) rpn_filter_fused_kernel(const RpnFusedParams p) {
  unsigned bar_gen = 0;
  extern __shared__ __align__(16) unsigned char fu_smem[];
  __shared__ int s_hist[kSelBins];
  __shared__ int s_scan[32];
  __shared__ int s_res[3];
  __shared__ int s_n[kFuMaxGroups], s_start[kFuMaxGroups + 1];
  __shared__ uint32_t s_cur[32];
  __shared__ int s_cnt[32];
  __shared__ int s_base;
  const int tid = threadIdx.x;
  const unsigned nblocks = gridDim.x;
  const FusedWs& ws = p.ws;

  // ---- P0: exact per-level top-k by radix select (rpn_function.py:447-456), one digit per grid barrier ----
  int l = 0;
  while (l + 1 < p.num_levels && (int)blockIdx.x >= p.cta_begin[l + 1]) ++l;
  const int x = blockIdx.x - p.cta_begin[l], nx = p.cta_begin[l + 1] - p.cta_begin[l];
  const int begin = p.lv.start[l], end = p.lv.start[l + 1];
  SelState st;
  st.prefix = 0; st.k_rem = p.topk.k[l]; st.done = 0;
  for (int pass = 0; pass <= kSelPasses; ++pass) {
    if (begin < end) {  // CTA-uniform
      if (pass > 0) st = sel_resolve_1024(st, p.hist + ((size_t)(pass - 1) * p.num_levels + l) * kSelBins, pass - 1, s_scan, s_res);
      if (pass < kSelPasses) {
        if (!st.done) {
          for (int i = tid; i < kSelBins; i += kFuThreads) s_hist[i] = 0;
          __syncthreads();
          const int shift = kSelShift[pass], dmask = (1 << kSelBits[pass]) - 1;
          for (int i = begin + x * kFuThreads + tid; i < end; i += nx * kFuThreads) {
            const uint64_t key = rpn_rank_key(p.objectness[i], (uint32_t)i);
            if (pass == 0 || (key >> (shift + kSelBits[pass])) == st.prefix) atomicAdd(&s_hist[(int)(key >> shift) & dmask], 1);
          }
          __syncthreads();
          int* gh = p.hist + ((size_t)pass * p.num_levels + l) * kSelBins;
          for (int i = tid; i < kSelBins; i += kFuThreads)
            if (s_hist[i]) atomicAdd(&gh[i], s_hist[i]);
        }
      } else {
        const uint64_t kth = st.prefix;  // all 52 bits resolved
        uint64_t* dst = ws.lkey + (size_t)l * kFuCap;
        for (int i = begin + x * kFuThreads + tid; i < end; i += nx * kFuThreads) {
          const float logit = p.objectness[i];
          if (rpn_rank_key(logit, (uint32_t)i) <= kth) dst[atomicAdd(&ws.gcount[l], 1)] = make_cand_key((uint32_t)l, logit, (uint32_t)i);
        }
      }
    }
    grid_barrier(ws.bar, nblocks, bar_gen);
  }

  // ---- P1: a CTA per level: order the level's candidates (:455 topk order), clip, drop small / low-score boxes (:502-517) ----
  for (int g = blockIdx.x; g < p.num_levels; g += nblocks) {
    const int n = p.topk.k[g];  // exactly this many keys were emitted
    uint64_t* s_key = reinterpret_cast<uint64_t*>(fu_smem);
    const int len = pow2_at_least(n);
    if (tid < len) s_key[tid] = tid < n ? __ldcg(ws.lkey + (size_t)g * kFuCap + tid) : kPadKey;
    cta_bitonic(s_key, len);
    bool ok = false;
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    float sc = 0.f;
    if (tid < n) {
      const uint64_t key = s_key[tid];
      b = clip_box(p.proposals[cand_row(key)], p.img_w, p.img_h);                                     // :509
      sc = sigmoidf_ref(float_from_desc_bits(cand_score_bits(key)));                                   // :502
      ok = (b.z - b.x >= p.min_size) && (b.w - b.y >= p.min_size) && (sc >= p.score_thresh);           // :512, :517
    }
    int total;
    const int slot = cta_exclusive_scan(ok ? 1 : 0, s_scan, &total);
    if (ok) {
      const size_t at = (size_t)g * kFuCap + slot;
      ws.sbox[at] = b;
      ws.sscore[at] = sc;
      ws.spos[at] = (uint32_t)(g * kFuCap + slot);  // (level, rank in level): the order of the concatenated filtered lists
    }
    fused_max_coord(ok, b, ws.gmax);
    if (tid == 0) ws.gcount[g] = total;
    __syncthreads();
  }
  grid_barrier(ws.bar, nblocks, bar_gen);

  // ---- P2 .. P4: batched_nms (:520) as per-level NMS, keep[:post_nms_top_n] (:523) in result order ----
  fused_mask_phase(ws, p.num_levels, 0, p.thr, fu_smem, s_n, s_start, s_scan);
  grid_barrier(ws.bar, nblocks, bar_gen);
  fused_scan_phase(ws, p.num_levels, p.post_top_n, s_cur, s_cnt, &s_base);
  grid_barrier(ws.bar, nblocks, bar_gen);
  fused_merge_phase(ws, p.post_top_n, 0, fu_smem, p.out_boxes, p.out_scores, nullptr, p.out_count);
}

// ======================================================================================================================
// isb_box_postprocess in one launch
// ======================================================================================================================
struct BoxFusedParams {
  BoxPostParams bp;
  float thr;
  int max_keep;
  FusedWs ws;
  float4* out_boxes;
  float* out_scores;
  int64_t* out_labels;
  int32_t* out_count;
};

__global__ void __launch_bounds__(kFuThreads, 1) boxpost_fused_kernel(const BoxFusedParams q) {
  unsigned bar_gen = 0;
  extern __shared__ __align__(16) unsigned char fu_smem[];
  __shared__ int s_scan[32];
  __shared__ int s_n[kFuMaxGroups], s_start[kFuMaxGroups + 1];
  __shared__ uint32_t s_cur[32];
  __shared__ int s_cnt[32];
  __shared__ int s_base;
  const int tid = threadIdx.x, lane = tid & 31;
  const unsigned nblocks = gridDim.x;
  const FusedWs& ws = q.ws;
  const BoxPostParams& p = q.bp;
  const int groups = p.C - 1;  // group g = foreground class g + 1

  // ---- P0: a warp per proposal row: softmax (:401), decode (:398), clip (:404), score / size filters (:437-442);
  //          every survivor joins its class's list with the key (score, row * C + class) -- at most one per row and class ----
  const int rows = boxpost_rows(p);
  for (int r = blockIdx.x * kFuWarps + (tid >> 5); r < rows; r += nblocks * kFuWarps) {
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
    for (int c0 = 1; c0 < p.C; c0 += 32) {  // background class 0 dropped (roi_head.py:423-425)
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
      if (ok) {
        const int slot = atomicAdd(&ws.gcount[c - 1], 1);  // < R <= cap: a row adds at most one candidate to a class
        const size_t at = (size_t)(c - 1) * kFuCap + slot;
        // ~score (32) | list position row * C + class (20) | slot (10)
        ws.lkey[at] = ((uint64_t)float_desc_bits(sc) << 30) | ((uint64_t)(r * p.C + c) << 10) | (uint64_t)slot;
        ws.lbox[at] = b;
      }
      fused_max_coord(ok, b, ws.gmax);
    }
  }
  grid_barrier(ws.bar, nblocks, bar_gen);

  // ---- P1: a CTA per class: candidates in score order (ties by list position) ----
  for (int g = blockIdx.x; g < groups; g += nblocks) {
    const int n = __ldcg(ws.gcount + g);
    if (n <= 0) continue;  // CTA-uniform
    uint64_t* s_key = reinterpret_cast<uint64_t*>(fu_smem);
    const int len = pow2_at_least(n);
    if (tid < len) s_key[tid] = tid < n ? __ldcg(ws.lkey + (size_t)g * kFuCap + tid) : kPadKey;
    cta_bitonic(s_key, len);
    if (tid < n) {
      const uint64_t key = s_key[tid];
      const size_t at = (size_t)g * kFuCap + tid;
      ws.sbox[at] = __ldcg(ws.lbox + (size_t)g * kFuCap + (int)(key & 1023u));
      ws.sscore[at] = float_from_desc_bits((uint32_t)(key >> 30));
      ws.spos[at] = (uint32_t)(key >> 10) & 0xfffffu;
    }
    __syncthreads();
  }
  grid_barrier(ws.bar, nblocks, bar_gen);

  // ---- P2 .. P4: batched_nms over the labels (:445) as per-class NMS, keep[:detections_per_img] (:448) ----
  fused_mask_phase(ws, groups, 1, q.thr, fu_smem, s_n, s_start, s_scan);
  grid_barrier(ws.bar, nblocks, bar_gen);
  fused_scan_phase(ws, groups, q.max_keep, s_cur, s_cnt, &s_base);
  grid_barrier(ws.bar, nblocks, bar_gen);
  fused_merge_phase(ws, q.max_keep, 1, fu_smem, q.out_boxes, q.out_scores, q.out_labels, q.out_count);
}

}  // namespace isb
