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
constexpr size_t kNmsSmem = (size_t)kSmemCand * (8 + 8 + 16);  // keys, kept keys, boxes

// Coarse descending bin of a detection key's score, monotone along the key order (best scores -> bin 0): linear in
// [0, 1] (class probabilities: ~uniform resolution where it matters), everything above 1 in bin 0, everything below 0
// in the last bin.  Only used to pre-select a superset of the top max_det keys, so resolution affects speed, not results.
__device__ __forceinline__ int score_bin_desc(uint64_t det_key) {
  const float s = det_score(det_key);
  if (!(s == s)) return (__float_as_uint(s) >> 31) ? 2047 : 0;  // -NaN sorts last, +NaN first
  const float t = fminf(fmaxf(s * 2048.0f, 0.0f), 2047.0f);
  return 2047 - (int)t;
}

// Warp-aggregated slot allocation: every lane of the warp calls it; lanes with `want` get consecutive slots from one
// shared-memory atomicAdd per warp (hundreds of single-address atomics otherwise serialise for microseconds).
__device__ __forceinline__ int warp_alloc_slot(bool want, int* counter) {
  const unsigned bal = __ballot_sync(0xffffffffu, want);
  const int lane = threadIdx.x & 31;
  int base = 0;
  if (lane == 0 && bal) base = atomicAdd(counter, __popc(bal));
  base = __shfl_sync(0xffffffffu, base, 0);
  return base + __popc(bal & ((1u << lane) - 1u));
}

// Compare-exchange of register-resident keys across the lanes of a warp: element index i, partner i ^ stride.
__device__ __forceinline__ void cmpx_lane(uint64_t& v, int i, int stride, int size) {
  const uint64_t o = __shfl_xor_sync(0xffffffffu, v, stride);
  const bool lower = (i & stride) == 0;
  const bool asc = (i & size) == 0;  // size > stride, so both partners agree
  const bool take_min = lower == asc;
  v = ((v < o) == take_min) ? v : o;
}

// Strides 32, 16, .., 1 of merge level `size` (>= 64) for one 64-key block, in registers.
__device__ __forceinline__ void bitonic_tail64(uint64_t* k, int base, int size, int lane) {
  const int ia = base + lane, ib = ia + 32;
  uint64_t a = k[ia], b = k[ib];
  if ((a > b) == ((ia & size) == 0)) { const uint64_t t = a; a = b; b = t; }
#pragma unroll
  for (int stride = 16; stride > 0; stride >>= 1) {
    cmpx_lane(a, ia, stride, size);
    cmpx_lane(b, ib, stride, size);
  }
  k[ia] = a;
  k[ib] = b;
}

// Ascending bitonic sort of n keys (n a power of two >= 64) by the whole CTA.  Every compare-exchange
// with stride <= 32 runs on register-resident keys through warp shuffles (a warp owns a 64-key block);
// only the strides >= 64 go through memory with a CTA barrier: 18 barriers instead of 55 at n = 1024.
__device__ void bitonic_sort_u64(uint64_t* k, int n) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  __syncthreads();
  for (int blk = warp; blk < (n >> 6); blk += nwarps) {  // sizes 2..64 without touching memory in between
    const int base = blk << 6, ia = base + lane, ib = ia + 32;
    uint64_t a = k[ia], b = k[ib];
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        cmpx_lane(a, ia, stride, size);
        cmpx_lane(b, ib, stride, size);
      }
    }
    if ((a > b) == ((ia & 64) == 0)) { const uint64_t t = a; a = b; b = t; }
#pragma unroll
    for (int stride = 16; stride > 0; stride >>= 1) {
      cmpx_lane(a, ia, stride, 64);
      cmpx_lane(b, ib, stride, 64);
    }
    k[ia] = a;
    k[ib] = b;
  }
  for (int size = 128; size <= n; size <<= 1) {
    for (int stride = size >> 1; stride >= 64; stride >>= 1) {
      __syncthreads();
      for (int t = threadIdx.x; t < (n >> 1); t += blockDim.x) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool asc = (lo & size) == 0;
        const uint64_t a = k[lo], b = k[hi];
        if ((a > b) == asc) {
          k[lo] = b;
          k[hi] = a;
        }
      }
    }
    __syncthreads();
    for (int blk = warp; blk < (n >> 6); blk += nwarps) bitonic_tail64(k, blk << 6, size, lane);
  }
  __syncthreads();
}

__device__ __forceinline__ float4 load_box(const NmsParams& p, int b, uint32_t row) {
  if (p.boxes) return p.boxes[(size_t)b * p.rows + row];
  float4 pr = p.pred_box[(size_t)b * p.rows + row];
  float4 an = p.anchors[row];
  return decode_box(pr, an, 0.1f, 0.2f);
}

// Outputs of one image (step 5 of fast_nms_kernel, shared with the wide path): boxes (rescaled / clamped when asked),
// scores, class ids, coefficient rows, and for the mask kernels the detection records + sigmoid-map offsets.
// Whole CTA (kNmsThreads); dkeys = sorted kept keys, m <= max_det <= kNmsThreads of them.
__device__ __forceinline__ void fnms_emit(const NmsParams& p, int b, const uint64_t* dkeys, int m, float* __restrict__ det_coef_ws,
                                          DetRec* __restrict__ det_rec_ws, int32_t* __restrict__ smap_total_ws, int ws_det_stride,
                                          int* s_scan) {
  const int tid = threadIdx.x;
  const float wf = (float)p.out_w, hf = (float)p.out_h;
  const float sy = (float)p.proto_h / hf, sx = (float)p.proto_w / wf;
  for (int d0 = 0; d0 < m; d0 += kNmsThreads) {  // max_det <= 1024: one trip
    const int d = d0 + tid;
    DetRec rc;
    int items = 0;
    if (d < m) {
      const uint64_t k = dkeys[d];
      const uint32_t row = det_row(k);
      float4 bx = load_box(p, b, row);
      if (p.correct) {
        bx = correct_box(bx, wf, hf);
        DetBounds db;
        crop_range(bx.x, bx.z, p.out_w, db.xlo, db.xhi);
        crop_range(bx.y, bx.w, p.out_h, db.ylo, db.yhi);
        int pa, pya, fw, fh;
        det_footprint(db, sy, sx, p.proto_h, p.proto_w, pa, pya, fw, fh);
        rc.xlo = db.xlo; rc.xhi = db.xhi; rc.ylo = db.ylo; rc.yhi = db.yhi;
        rc.pa = (int16_t)pa; rc.pya = (int16_t)pya; rc.fw = (int16_t)fw; rc.fh = (int16_t)fh;
        rc.j = d;
        items = fw * fh;
      }
      reinterpret_cast<float4*>(p.out_box)[(size_t)b * p.max_det + d] = bx;
      p.out_score[(size_t)b * p.max_det + d] = det_score(k);
      p.out_class[(size_t)b * p.max_det + d] = (int64_t)det_class(k);
    }
    if (p.correct) {  // offsets of the detections' sigmoid maps in the per-image scratch (mask kernels)
      int total;
      rc.soff = cta_exclusive_scan(items, s_scan, &total);
      if (d < m) det_rec_ws[(size_t)b * ws_det_stride + d] = rc;
      if (tid == 0) smap_total_ws[b] = total;  // max_det <= 1024: the single trip's total is the image's
    }
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
}

// counts[b] = m, and -- multi-GPU jobs with the fused count exchange attached -- the same count into the current ring slot of
// every rank over NVLink: one self-validating 64-bit word (epoch << 32 | count) per image and peer, written by the
// CTA's first thread.  Called by the whole CTA.  The word is the whole message (nothing else has to be visible before it),
// so the stores are relaxed: eight release stores at system scope would each wait for the CTA's earlier writes to drain.
__device__ __forceinline__ void publish_count(const NmsParams& p, int b, int m) {
  if (threadIdx.x == 0) {
    p.out_count[b] = m;
    const unsigned long long w = ((unsigned long long)p.peer_epoch << 32) | (unsigned)m;
#pragma unroll
    for (int r = 0; r < 8; ++r)  // static indices: the pointer table stays in the kernel's parameter bank
      if (r < p.n_peers)
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.peer_ring[r] + p.peer_index0 + b), "l"(w) : "memory");
  }
}

// utils_bbox.py:38-47 + :86 for one pair: does box i (higher ranked) suppress box j?  The division is skipped when the
// boxes are disjoint: 0 / union is +-0 for every finite non-zero union, and the caller established 0 <= nms_iou.
__device__ __forceinline__ bool fnms_suppresses(float4 bi, float4 bj, float aj, float thr) {
  const float dx = fmaxf(fminf(bi.z, bj.z) - fmaxf(bi.x, bj.x), 0.0f);
  const float dy = fmaxf(fminf(bi.w, bj.w) - fmaxf(bi.y, bj.y), 0.0f);
  const float inter = dx * dy;
  const float uni = (box_area(bi) + aj) - inter;
  if (inter == 0.0f && fabsf(uni) > 0.0f) return false;  // non-zero, not NaN (an infinite union gives 0 too)
  return !(div_rn_guarded(inter, uni) <= thr);
}

__global__ void __launch_bounds__(kNmsThreads, 1)
fast_nms_kernel(NmsParams p, const int32_t* __restrict__ cand_count, uint64_t* __restrict__ g_keys,
                uint64_t* __restrict__ g_det, float4* __restrict__ g_box, int64_t cap, int64_t stride,
                float* __restrict__ det_coef_ws, DetRec* __restrict__ det_rec_ws, int32_t* __restrict__ smap_total_ws,
                int ws_det_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_kept;
  __shared__ int s_scan[32];
  __shared__ int s_cls_start[1 << kClassBits];
  __shared__ int s_cut;
  __shared__ __align__(16) uint8_t s_sup[kSmemCand];  // sort-free walk: candidate is suppressed / outside its class window

  // the sigmoid-map kernel behind this one may be scheduled onto the idle SMs now (it waits before reading anything)
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");  // (launched as candidates_kernel's programmatic dependent in the fused path)
  const int b = blockIdx.x;
  const int tid = threadIdx.x;
  const int64_t total = cand_count[b];
  if (total > cap) {  // candidate buffer overflowed: report, never guess
    publish_count(p, b, -1);
    return;
  }
  const int n = (int)total;
  if (n == 0) {
    publish_count(p, b, 0);
    return;
  }
  const int np2 = max(64, next_pow2(n));
  const bool in_smem = np2 <= kSmemCand;
  uint64_t* keys = in_smem ? reinterpret_cast<uint64_t*>(smem_raw) : g_keys + (size_t)b * stride;
  uint64_t* dkeys = in_smem ? reinterpret_cast<uint64_t*>(smem_raw) + kSmemCand : g_det + (size_t)b * stride;
  float4* cbox = in_smem ? reinterpret_cast<float4*>(smem_raw + 2 * kSmemCand * sizeof(uint64_t))
                         : g_box + (size_t)b * stride;

  // ---- 1-3. class lists, boxes, Fast NMS ---------------------------------------------------
  const bool thr_ok = 0.0f <= p.nms_iou;  // column max of a triu'd matrix is >= 0 (:80-86)
  const int sub = tid & 7;
  const unsigned gmask = 0xffu << (tid & 24);  // the 8 lanes of this group
  if (tid == 0) { s_kept = 0; s_cut = 0; }
  bool done = false;  // CTA-uniform
  if (in_smem) {
    // No sort at all when the class lists are short (the usual case: ~1000 candidates over dozens of classes): a counting
    // sort BY CLASS only (shared-memory histogram, scan, scatter), then every candidate walks its class mates once and
    // both counts the ones that rank before it (its place in the class's sorted list, :61-64) and tests them as
    // suppressors (:75-86).  That is sum(class size^2) / 8 steps of an 8-lane group -- a few thousand warp instructions
    // where the bitonic network + class table + triangular walk spend ~60 k.  The order inside a class never has to be
    // materialised: kept candidates are re-sorted by score below anyway.  Images whose classes are too long for the
    // quadratic walk (one class with thousands of candidates) take the bitonic sort.
    // pointers derived straight from the shared array, so that the walks below compile to LDS / STS (the generic
    // `keys` / `cbox` of the spill path make every access a generic load with its address arithmetic)
    uint64_t* const skeys = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* const raw = skeys + kSmemCand;  // raw keys first live in the second buffer
    float4* const sbox = reinterpret_cast<float4*>(smem_raw + 2 * kSmemCand * sizeof(uint64_t));
    const uint64_t* src = g_keys + (size_t)b * stride;
    for (int i = tid; i < n; i += kNmsThreads) raw[i] = src[i];
    for (int c = tid; c < (1 << kClassBits); c += kNmsThreads) s_cls_start[c] = 0;
    for (int i = tid; i < (n + 3) / 4; i += kNmsThreads) reinterpret_cast<uint32_t*>(s_sup)[i] = 0u;
    __syncthreads();
    for (int i = tid; i < n; i += kNmsThreads) atomicAdd(&s_cls_start[cand_class(raw[i])], 1);
    __syncthreads();
    constexpr int kPer = (1 << kClassBits) / kNmsThreads;
    int cnt[kPer], sum = 0, sq = 0;
#pragma unroll
    for (int q = 0; q < kPer; ++q) {
      cnt[q] = s_cls_start[tid * kPer + q];
      sum += cnt[q];
      sq += cnt[q] * cnt[q];
    }
    sq = __reduce_add_sync(0xffffffffu, sq);
    if ((tid & 31) == 0 && sq) atomicAdd(&s_cut, sq);
    int tot;
    int run = cta_exclusive_scan(sum, s_scan, &tot);  // barriers inside: s_cut is complete afterwards
#pragma unroll
    for (int q = 0; q < kPer; ++q) {
      s_cls_start[tid * kPer + q] = run;
      run += cnt[q];
    }
    __syncthreads();
    if (s_cut <= 64 * np2) {
      for (int i = tid; i < n; i += kNmsThreads) {
        const uint64_t k = raw[i];
        const int pos = atomicAdd(&s_cls_start[cand_class(k)], 1);  // afterwards s_cls_start[c] = END of class c
        skeys[pos] = k;
        sbox[pos] = load_box(p, b, cand_row(k));
      }
      __syncthreads();
      // A list no longer than top_k lies wholly inside the window (:63-64), so ranks are not needed and every unordered
      // pair is looked at ONCE (i before j in the list's arbitrary order): whichever of the two ranks lower is flagged when
      // they overlap too much -- a suppressed box still suppresses (Fast NMS), so nothing stops early.  Longer lists walk
      // all mates of j, count the ones that rank before it and flag j alone.
      for (int j0 = 0; j0 < n; j0 += kNmsThreads / 8) {
        const int j = j0 + (tid >> 3);
        bool bad = false;
        if (j < n) {
          const uint64_t kj = skeys[j];
          const uint32_t c = cand_class(kj);
          const int lo = c ? s_cls_start[c - 1] : 0, hi = s_cls_start[c];
          const float4 bj = sbox[j];
          const float aj = box_area(bj);
          if (hi - lo <= p.top_k) {
#pragma unroll 1
            for (int i = lo + sub; i < j; i += 8) {
              const float4 bi = sbox[i];
              if (fnms_suppresses(bi, bj, aj, p.nms_iou)) {  // symmetric in the two boxes
                if (skeys[i] < kj) bad = true; else s_sup[i] = 1;
              }
            }
          } else {
            int before = 0;
#pragma unroll 1
            for (int i = lo + sub; i < hi; i += 8) {
              const bool higher = skeys[i] < kj;  // i ranks before j (score desc, row asc); false for i == j
              before += higher;
              if (higher && fnms_suppresses(sbox[i], bj, aj, p.nms_iou)) bad = true;
            }
            before += __shfl_xor_sync(gmask, before, 1);
            before += __shfl_xor_sync(gmask, before, 2);
            before += __shfl_xor_sync(gmask, before, 4);
            if (before >= p.top_k) bad = true;
          }
          if (bad) s_sup[j] = 1;
        }
      }
      __syncthreads();
      for (int j0 = 0; j0 < n; j0 += kNmsThreads) {
        const int j = j0 + tid;
        const bool emit = thr_ok && j < n && !s_sup[j];
        const int slot = warp_alloc_slot(emit, &s_kept);
        if (emit) raw[slot] = det_key_from_cand(skeys[j]);  // the raw keys are no longer needed (raw == dkeys)
      }
      done = true;
    } else {
      for (int i = tid; i < np2; i += kNmsThreads) keys[i] = i < n ? raw[i] : kPadKey;
      bitonic_sort_u64(keys, np2);
    }
  } else {
    for (int i = n + tid; i < np2; i += kNmsThreads) keys[i] = kPadKey;
    bitonic_sort_u64(keys, np2);
  }
  if (!done) {
    // sorted keys (class asc, score desc, row asc): boxes, first position of every class, triangular walk
    for (int i = tid; i < n; i += kNmsThreads) cbox[i] = load_box(p, b, cand_row(keys[i]));
    __syncthreads();
    for (int i = tid; i < n; i += kNmsThreads) {
      const uint32_t c = cand_class(keys[i]);
      if (i == 0 || cand_class(keys[i - 1]) != c) s_cls_start[c] = i;
    }
    __syncthreads();
    // EIGHT lanes per candidate j, striding over its higher-ranked class mates: a class with a long list does not
    // leave one thread walking it alone while the CTA waits, and the per-candidate set-up is shared by 4 candidates
    // per warp instruction.
    for (int j0 = 0; j0 < n; j0 += kNmsThreads / 8) {
      const int j = j0 + (tid >> 3);
      bool bad = true;
      uint64_t kj = 0;
      if (j < n) {
        kj = keys[j];
        const int start = s_cls_start[cand_class(kj)];
        if (thr_ok && (j - start) < p.top_k) {
          bad = false;
          const float4 bj = cbox[j];
          const float aj = box_area(bj);
          for (int i = start + sub; i < j; i += 8) {
            if (fnms_suppresses(cbox[i], bj, aj, p.nms_iou)) { bad = true; break; }
          }
        }
      }
      const unsigned any_bad = __ballot_sync(0xffffffffu, bad) & gmask;
      const bool emit = !any_bad && sub == 0;
      const int slot = warp_alloc_slot(emit, &s_kept);
      if (emit) dkeys[slot] = det_key_from_cand(kj);
    }
  }
  __syncthreads();
  // ---- 4. global order: score desc, class asc, row asc; keep max_det --------------------
  // Only the best max_det kept candidates are wanted (:94-95): histogram a coarse descending bin of the score
  // (score_bin_desc), find the bin where the running count reaches max_det, and sort just the candidates up to that bin (a
  // superset of the answer, typically ~1.2 x max_det instead of every kept candidate).
  int kept = s_kept;
  if (kept > 2 * p.max_det && kept > 128 && kept <= kSmemCand) {
    int* s_hist = s_cls_start;  // 2048 bins (the class table is no longer needed); kClassBits >= 11
    for (int i = tid; i < 2048; i += kNmsThreads) s_hist[i] = 0;
    if (tid == 0) s_kept = 0;
    __syncthreads();
    for (int i = tid; i < kept; i += kNmsThreads) atomicAdd(&s_hist[score_bin_desc(dkeys[i])], 1);
    __syncthreads();
    const int h0 = s_hist[2 * tid], h1 = s_hist[2 * tid + 1];
    int total;
    const int excl = cta_exclusive_scan(h0 + h1, s_scan, &total);
    // bins strictly before `cut` hold < max_det candidates in total; every bin <= cut is taken
    if (excl < p.max_det && excl + h0 + h1 >= p.max_det) s_cut = (excl + h0 >= p.max_det) ? 2 * tid : 2 * tid + 1;
    __syncthreads();
    const int cut = s_cut;
    // compact in place: a candidate moves to a slot <= its own only after every thread has read its key
    uint64_t mine[(kSmemCand + kNmsThreads - 1) / kNmsThreads];
#pragma unroll
    for (int q = 0; q < (kSmemCand + kNmsThreads - 1) / kNmsThreads; ++q)
      if (q * kNmsThreads + tid < kept) mine[q] = dkeys[q * kNmsThreads + tid];
    __syncthreads();
#pragma unroll
    for (int q = 0; q < (kSmemCand + kNmsThreads - 1) / kNmsThreads; ++q)
    {
      const bool take = q * kNmsThreads + tid < kept && score_bin_desc(mine[q]) <= cut;
      const int slot = warp_alloc_slot(take, &s_kept);
      if (take) dkeys[slot] = mine[q];
    }
    __syncthreads();
    kept = s_kept;
  }
  const int kp2 = max(64, next_pow2(kept));
  for (int i = kept + tid; i < kp2; i += kNmsThreads) dkeys[i] = kPadKey;
  bitonic_sort_u64(dkeys, kp2);
  const int m = min(kept, p.max_det);

  // ---- 5. outputs -----------------------------------------------------------------------
  fnms_emit(p, b, dkeys, m, det_coef_ws, det_rec_ws, smap_total_ws, ws_det_stride, s_scan);
  publish_count(p, b, m);
}


// ---------------------------------------------------------------------------------------------------------------
// Wide path for ONE image with many candidates (BASELINE configs[4], VerSeg: 5000 boxes of one class, top_k 5000,
// max_det 1024).  fast_nms_kernel gives an image to a single CTA, which for n > kSmemCand sorts in global memory and
// walks n^2 / 2 box pairs alone (4.9 ms for n = 5000).  Here the same steps are spread over the GPU: multi-CTA bitonic
// sort of the keys, boxes + class starts, pair tests with 32 candidates per CTA, sort of the kept keys, emit.
// ---------------------------------------------------------------------------------------------------------------
__global__ void fnms_fill_kernel(uint64_t* __restrict__ keys, int from, int to, int32_t* __restrict__ zero) {
  pdl_enter();
  for (int i = from + blockIdx.x * blockDim.x + threadIdx.x; i < to; i += gridDim.x * blockDim.x) keys[i] = kPadKey;
  if (zero && blockIdx.x == 0 && threadIdx.x == 0) *zero = 0;
}

__global__ void fnms_prepare_kernel(NmsParams p, const uint64_t* __restrict__ keys, int n, float4* __restrict__ cbox,
                                    int32_t* __restrict__ cls_start) {
  pdl_enter();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint64_t k = keys[i];
    cbox[i] = load_box(p, 0, cand_row(k));
    const uint32_t c = cand_class(k);
    if (i == 0 || cand_class(keys[i - 1]) != c) cls_start[c] = i;
  }
}

constexpr int kWideThreads = 256;  // 8 candidates per CTA, one warp each

// A WARP per candidate j, four independent box loads in flight per lane: with thousands of predecessors per candidate
// (one class, VerSeg) an 8-lane group with one load in flight spent ~100 us waiting on L2.
__global__ void __launch_bounds__(kWideThreads)
fnms_pairs_kernel(NmsParams p, const uint64_t* __restrict__ keys, const float4* __restrict__ cbox,
                  const int32_t* __restrict__ cls_start, int n, uint64_t* __restrict__ dkeys, int32_t* __restrict__ kept) {
  pdl_enter();
  const int lane = threadIdx.x & 31;
  const bool thr_ok = 0.0f <= p.nms_iou;
  // the longest lists (highest j of a class) first: they decide when the kernel ends
  const int j = n - 1 - (blockIdx.x * (kWideThreads / 32) + (threadIdx.x >> 5));
  if (j < 0) return;  // warp-uniform
  const uint64_t kj = keys[j];
  const int start = cls_start[cand_class(kj)];
  bool bad = !(thr_ok && (j - start) < p.top_k);
  if (!bad) {
    const float4 bj = cbox[j];
    const float aj = box_area(bj);
    for (int i0 = start; i0 < j; i0 += 128) {
      float4 bb[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + 32 * u + lane;
        bb[u] = cbox[i < j ? i : j];
      }
      bool hit = false;
#pragma unroll
      for (int u = 0; u < 4; ++u) hit |= (i0 + 32 * u + lane < j) && fnms_suppresses(bb[u], bj, aj, p.nms_iou);
      if (__any_sync(0xffffffffu, hit)) { bad = true; break; }
    }
  }
  if (lane == 0 && !bad) dkeys[atomicAdd(kept, 1)] = det_key_from_cand(kj);
}

__global__ void __launch_bounds__(kNmsThreads, 1)
fnms_emit_kernel(NmsParams p, const uint64_t* __restrict__ dkeys, const int32_t* __restrict__ kept, float* __restrict__ det_coef_ws,
                 DetRec* __restrict__ det_rec_ws, int32_t* __restrict__ smap_total_ws, int ws_det_stride) {
  __shared__ int s_scan[32];
  pdl_enter();
  const int m = min(*kept, p.max_det);
  fnms_emit(p, 0, dkeys, m, det_coef_ws, det_rec_ws, smap_total_ws, ws_det_stride, s_scan);
  if (threadIdx.x == 0) p.out_count[0] = m;
}

int launch_fast_nms_wide(isb_handle* h, const NmsParams& p, int n, cudaStream_t s) {
  if (p.batch != 1) ISB_FAIL(h, ISB_ERR_BAD_ARG, "wide Fast NMS handles one image");
  const size_t b0 = (size_t)h->ws_b0, cap2 = (size_t)h->cand_cap_pow2;
  const int np2 = max(64, next_pow2(n));
  if ((size_t)np2 > cap2) ISB_FAIL(h, ISB_ERR_LIMIT, "wide Fast NMS: %d candidates exceed the key buffer", n);
  uint64_t* keys = h->cand_keys + b0 * cap2;
  uint64_t* dkeys = h->det_keys + b0 * cap2;
  float4* cbox = h->cand_box + b0 * cap2;
  if (!h->fnms_cls) ISB_CUDA(h, cudaMalloc((void**)&h->fnms_cls, sizeof(int32_t) * ((1 << kClassBits) + 1)));
  int32_t* cls_start = h->fnms_cls;
  int32_t* kept = h->fnms_cls + (1 << kClassBits);
  int rc;
  float* coef_ws = h->det_coef + b0 * h->lim.max_det * ISB_NUM_COEF;
  DetRec* rec_ws = h->det_rec + b0 * h->lim.max_det;
  const int pair_grid = (n + kWideThreads / 32 - 1) / (kWideThreads / 32);
  if (n <= kRankSortMax) {
    // counting rank sorts (multi-CTA, two launches each); the box buffer doubles as their int32 rank scratch while it
    // holds no boxes.  keys -> sorted in dkeys -> boxes -> kept keys back into `keys` -> sorted in dkeys -> emit.
    int32_t* rank = reinterpret_cast<int32_t*>(cbox);
    if ((rc = launch_rank_sort_u64(h, keys, n, nullptr, rank, dkeys, s)) != ISB_OK) return rc;
    // every kernel from here on is launched as the programmatic dependent of the one before it (pdl_enter): the chain is
    // eight short kernels, and the launch latency between them was a sixth of its 76 us
    cudaError_t le;
    le = launch_chained(fnms_prepare_kernel, dim3((n + 255) / 256), dim3(256), 0, s, true, p, (const uint64_t*)dkeys, n, cbox, cls_start);
    if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "launch fnms_prepare_kernel: %s", cudaGetErrorString(le));
    le = launch_chained(fnms_fill_kernel, dim3(1), dim3(32), 0, s, true, keys, 0, 0, kept);  // kept = 0
    if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "launch fnms_fill_kernel: %s", cudaGetErrorString(le));
    le = launch_chained(fnms_pairs_kernel, dim3(pair_grid), dim3(kWideThreads), 0, s, true, p, (const uint64_t*)dkeys,
                        (const float4*)cbox, (const int32_t*)cls_start, n, keys, kept);
    if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "launch fnms_pairs_kernel: %s", cudaGetErrorString(le));
    h->launches += 3;
    if ((rc = launch_rank_sort_u64(h, keys, n, kept, rank, dkeys, s)) != ISB_OK) return rc;
    le = launch_chained(fnms_emit_kernel, dim3(1), dim3(kNmsThreads), 0, s, true, p, (const uint64_t*)dkeys, (const int32_t*)kept,
                        coef_ws, rec_ws, h->smap_total + b0, (int)h->lim.max_det);
    if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "launch fnms_emit_kernel: %s", cudaGetErrorString(le));
    h->launches++;
    return ISB_OK;
  }
  fnms_fill_kernel<<<8, 256, 0, s>>>(keys, n, np2, nullptr);
  ISB_LAUNCH_CHECK(h, "fnms_fill_kernel");
  if ((rc = launch_sort_u64(h, keys, np2, s)) != ISB_OK) return rc;
  fnms_prepare_kernel<<<(n + 255) / 256, 256, 0, s>>>(p, keys, n, cbox, cls_start);
  ISB_LAUNCH_CHECK(h, "fnms_prepare_kernel");
  fnms_fill_kernel<<<(np2 + 255) / 256, 256, 0, s>>>(dkeys, 0, np2, kept);
  ISB_LAUNCH_CHECK(h, "fnms_fill_kernel");
  fnms_pairs_kernel<<<pair_grid, kWideThreads, 0, s>>>(p, keys, cbox, cls_start, n, dkeys, kept);
  ISB_LAUNCH_CHECK(h, "fnms_pairs_kernel");
  if ((rc = launch_sort_u64(h, dkeys, np2, s)) != ISB_OK) return rc;
  fnms_emit_kernel<<<1, kNmsThreads, 0, s>>>(p, dkeys, kept, coef_ws, rec_ws, h->smap_total + b0, h->lim.max_det);
  ISB_LAUNCH_CHECK(h, "fnms_emit_kernel");
  return ISB_OK;
}

int fast_nms_smem_candidates() { return kSmemCand; }

int configure_fast_nms(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(fast_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kNmsSmem));
  return ISB_OK;
}

int launch_fast_nms(isb_handle* h, const NmsParams& p, cudaStream_t s) {
  const size_t b0 = (size_t)h->ws_b0, cap2 = (size_t)h->cand_cap_pow2;
  cudaLaunchConfig_t lc;
  memset(&lc, 0, sizeof(lc));
  lc.gridDim = dim3((unsigned)p.batch);
  lc.blockDim = dim3(kNmsThreads);
  lc.dynamicSmemBytes = kNmsSmem;
  lc.stream = s;
  cudaLaunchAttribute attr;
  memset(&attr, 0, sizeof(attr));
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  static const bool no_pdl = getenv("ISB_NO_PDL") != nullptr;
  lc.attrs = &attr;
  lc.numAttrs = (h->nms_follows_cand && !no_pdl) ? 1 : 0;
  const cudaError_t le = cudaLaunchKernelEx(
      &lc, fast_nms_kernel, p, (const int32_t*)(h->cand_count + b0), h->cand_keys + b0 * cap2, h->det_keys + b0 * cap2,
      h->cand_box + b0 * cap2, (int64_t)h->lim.cand_capacity, (int64_t)h->cand_cap_pow2,
      h->det_coef + b0 * h->lim.max_det * ISB_NUM_COEF, h->det_rec + b0 * h->lim.max_det, h->smap_total + b0, (int)h->lim.max_det);
  if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "fast_nms_kernel launch: %s", cudaGetErrorString(le));
  h->launches++;
  return ISB_OK;
}

}  // namespace isb
