// K1 -- score filter + candidate emission.
//
// Replaces utils_bbox.py:193-204 (drop background column, row max over classes, > confidence,
// boolean gather) and the "transpose + per-class sort input" of :55-61.  Instead of gathering
// n x (C-1) score rows, every (row, class) entry whose score passes the threshold becomes one
// 64-bit key (class | ~score | row); see DESIGN.md "Exact shortcut" for why entries <= confidence
// can never change the result.
//
// HBM-read-bound: the score tensor [B, A, C] is read exactly once (6.2 MB / image at COCO 550) as a flat
// stream of 16-byte vectors, four independent loads in flight per thread.  Passing entries are rare (about
// 1 in 1500 at confidence 0.05) apart from the background column, so the common path is
// load + 4 compares; only a vector with a passing entry pays for the (row, class) index arithmetic and one
// atomicAdd on the image's candidate counter.  The order of the emitted keys is irrelevant: the NMS kernel
// sorts them, and keys are unique.
#include "isb_common.cuh"

namespace isb {

constexpr int kCandThreads = 256;
constexpr int kCandUnroll = 4;

constexpr int kCandStage = 1024;  // keys a CTA parks in shared memory before its single reservation

struct CandStage {
  uint64_t keys[kCandStage];
  int n;
  int base;
};

// A hit must not make its warp wait for a global atomic's round trip (1-2 us when a whole image hammers one
// counter): with ~450 hits per SM that wait, not bandwidth, set the kernel's duration.  Hits are parked in the
// CTA's shared memory and the CTA reserves its range of the image's buffer once, at the end.
__device__ __forceinline__ void cand_emit(const CandParams& p, int b, uint32_t e, float s, int32_t* __restrict__ cand_count,
                                          uint64_t* __restrict__ dst, int64_t cap, CandStage& st) {
  const uint32_t row = e / (uint32_t)p.ncols;
  const int c = (int)(e - row * (uint32_t)p.ncols);
  if (c < p.col0) return;
  if (p.use_threshold) {
    const uint64_t key = make_cand_key((uint32_t)(c - p.col0), s, row);
    const int sp = atomicAdd(&st.n, 1);
    if (sp < kCandStage) {
      st.keys[sp] = key;
    } else {  // stage full (dense scores): fall back to the direct reservation
      const int pos = atomicAdd(&cand_count[b], 1);
      if (pos < cap) dst[pos] = key;  // past the capacity: dropped, NMS reports overflow
    }
  } else {
    dst[(size_t)row * (p.ncols - p.col0) + (c - p.col0)] = make_cand_key((uint32_t)(c - p.col0), s, row);
  }
}

__global__ void __launch_bounds__(kCandThreads)
candidates_kernel(CandParams p, int32_t* __restrict__ cand_count, uint64_t* __restrict__ cand_keys, int64_t cap,
                  int64_t key_stride) {
  asm volatile("griddepcontrol.launch_dependents;");  // fast_nms_kernel may be queued behind this grid's last wave
  __shared__ CandStage st;
  if (threadIdx.x == 0) st.n = 0;
  __syncthreads();
  const int b = blockIdx.y;
  const uint32_t per_img = (uint32_t)p.rows * (uint32_t)p.ncols;
  const float* img = p.scores + (size_t)b * per_img;
  uint64_t* dst = cand_keys + (size_t)b * key_stride;
  const float thr = p.threshold;
  const bool all = !p.use_threshold;
  // 16-byte aligned body of this image's flat score range; scalar head and tail
  const uint32_t head = min(per_img, (uint32_t)(((16 - (reinterpret_cast<uintptr_t>(img) & 15)) & 15) >> 2));
  const uint32_t nvec = (per_img - head) >> 2;
  const float4* body = reinterpret_cast<const float4*>(img + head);
  const uint32_t stride = gridDim.x * kCandThreads;
  uint32_t v = blockIdx.x * kCandThreads + threadIdx.x;
  // The dropped leading column (background, col0 == 1) passes the threshold on almost every row, so testing it
  // per hit would send nearly every warp down the emit path.  Track r = (flat index of the vector) mod ncols
  // incrementally instead and blank that one element before the compare.  Purely a shortcut: cand_emit rejects
  // the same elements.
  const uint32_t ncols = (uint32_t)p.ncols;
  const bool blank = p.col0 == 1 && ncols >= 4 && !all;
  const uint32_t step = (4u * stride) % ncols;
  uint32_t r = (head + 4u * v) % ncols;
  const float kBlank = -__int_as_float(0x7f800000);
  auto scan = [&](float4 x, uint32_t e, uint32_t rr) {
    if (blank) {
      const uint32_t k = rr == 0 ? 0u : ncols - rr;  // position of the column-0 element in this vector when < 4
      if (k == 0) x.x = kBlank;
      if (k == 1) x.y = kBlank;
      if (k == 2) x.z = kBlank;
      if (k == 3) x.w = kBlank;
    }
    if (all || x.x > thr || x.y > thr || x.z > thr || x.w > thr) {
      if (all || x.x > thr) cand_emit(p, b, e, x.x, cand_count, dst, cap, st);
      if (all || x.y > thr) cand_emit(p, b, e + 1, x.y, cand_count, dst, cap, st);
      if (all || x.z > thr) cand_emit(p, b, e + 2, x.z, cand_count, dst, cap, st);
      if (all || x.w > thr) cand_emit(p, b, e + 3, x.w, cand_count, dst, cap, st);
    }
  };
  for (; v + (kCandUnroll - 1) * stride < nvec; v += kCandUnroll * stride) {
    float4 x[kCandUnroll];
#pragma unroll
    for (int u = 0; u < kCandUnroll; ++u) x[u] = __ldcs(body + v + u * stride);  // streamed once: evict first
#pragma unroll
    for (int u = 0; u < kCandUnroll; ++u) {
      scan(x[u], head + 4 * (v + u * stride), r);
      r += step;
      if (r >= ncols) r -= ncols;
    }
  }
  for (; v < nvec; v += stride) {
    scan(__ldcs(body + v), head + 4 * v, r);
    r += step;
    if (r >= ncols) r -= ncols;
  }
  if (blockIdx.x == 0) {
    const uint32_t tail0 = head + 4 * nvec;
    for (uint32_t e = threadIdx.x; e < head; e += kCandThreads)
      if (all || img[e] > thr) cand_emit(p, b, e, img[e], cand_count, dst, cap, st);
    for (uint32_t e = tail0 + threadIdx.x; e < per_img; e += kCandThreads)
      if (all || img[e] > thr) cand_emit(p, b, e, img[e], cand_count, dst, cap, st);
    if (all && threadIdx.x == 0) cand_count[b] = p.rows * (p.ncols - p.col0);
  }
  if (all) return;
  __syncthreads();
  const int staged = min(st.n, kCandStage);
  if (staged == 0) return;
  if (threadIdx.x == 0) st.base = atomicAdd(&cand_count[b], staged);
  __syncthreads();
  for (int i = threadIdx.x; i < staged; i += kCandThreads) {
    const int64_t pos = (int64_t)st.base + i;
    if (pos < cap) dst[pos] = st.keys[i];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// K1L -- the same filter fed with RAW class logits: softmax over the C columns (nets/yolact.py:204,
// F.softmax(pred_classes, -1)) fused into the pass, so the conf tensor is never written and re-read.
//
// A row is C = 81 floats = 324 bytes: not a multiple of 16, so rows cannot be read as vectors, and a warp-per-row
// layout spends ~45 warp instructions per row on shuffles.  Instead tiles of 16 consecutive rows -- one contiguous 5 KB
// byte range -- are streamed into shared memory with one 1-D bulk (TMA) copy per tile.  A PAIR of warps shares a ring of
// three tile buffers (two tiles in flight while one is reduced): `full` mbarriers carry the copy's byte count, `empty`
// mbarriers (count 2) tell the pair's producer lane that both warps have left a buffer.  No CTA-wide barrier exists after
// start-up (a CTA-wide 128-row tile version spent 39 % of its time at the tile barrier waiting for warps whose rows had
// candidates).  Each warp reduces eight rows of the tile, FOUR lanes per row with 21 register-resident columns per lane:
// the first version (a warp per tile, a lane pair per row, 41 columns per lane) needed 122 registers and left 16 warps
// per SM to hide the latency of its dependent max / sum chains (ncu: 36 % of the issue slots used at 3.6 TB/s, while
// the tile pipeline alone runs at 5.9 TB/s); four lanes per row halve the chains and fit 24 warps (79 registers).
// Per row: the maximum, then -- only for rows whose foreground maximum is within log(threshold) of it, see lg_tile --
// the softmax denominator and the candidate test.  Hits wait in a per-warp list and leave with one reservation when the
// warp moves to another image (a pair's tiles are consecutive).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float exp2f_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

constexpr int kLgRows = 16;      // rows per tile = one bulk copy; a tile is shared by a pair of warps, eight rows each
constexpr int kLgThreads = 256;  // 8 warps
constexpr int kLgWarps = kLgThreads / 32;
constexpr int kLgPairs = kLgWarps / 2;
constexpr int kLgBufs = 2;       // tile buffers per pair
constexpr int kLgStage = 96;     // parked keys per warp
constexpr int kLgRegCols = 21;   // columns per lane kept in registers (four lanes per row: rows of up to 84 classes, COCO's 81)
constexpr int kLgWideRows = 2;   // up to this many candidate rows of a half tile are summed warp-wide, one after the other

struct LgWarpStage {
  uint64_t keys[kLgStage];
  int n;
  int pad;
};

struct LgEmit {
  LgWarpStage* sg;
  int32_t* cand_count;
  uint64_t* cand_keys;
  int64_t cap, key_stride;
};

__device__ __forceinline__ void lg_flush(const LgEmit& e, int img, int lane) {  // whole warp, converged
  __syncwarp();
  const int staged = min(e.sg->n, kLgStage);
  if (staged > 0) {
    int gbase = 0;
    if (lane == 0) gbase = atomicAdd(&e.cand_count[img], staged);
    gbase = __shfl_sync(0xffffffffu, gbase, 0);
    for (int i = lane; i < staged; i += 32)
      if ((int64_t)gbase + i < e.cap) e.cand_keys[(size_t)img * e.key_stride + gbase + i] = e.sg->keys[i];
  }
  __syncwarp();
  if (lane == 0) e.sg->n = 0;
  __syncwarp();
}

__device__ __forceinline__ void lg_emit_key(const LgEmit& e, int img, uint64_t key) {
  const int sp = atomicAdd(&e.sg->n, 1);
  if (sp < kLgStage) {
    e.sg->keys[sp] = key;
  } else {  // list full: reserve directly (slow, rare)
    const int pos = atomicAdd(&e.cand_count[img], 1);
    if (pos < e.cap) e.cand_keys[(size_t)img * e.key_stride + pos] = key;
  }
}

// One warp's half (rows 8 * half .. 8 * half + 7) of a tile of <= 16 rows, FOUR lanes per row (columns q, q + 4, ...).
// `tile` is either the shared-memory copy (the normal case; inlined separately so the loads compile to LDS) or, for the
// tensor's very last rows whose tail the aligned bulk copy could not cover, global memory.
// CT > 0: the class count as a compile-time constant (COCO's 81 and VOC's 21 columns are instantiated): the bounds test and
// select on every element of the register-resident quarter row fold away.
template <int CT>
__device__ __forceinline__ void lg_tile(const float* tile, const CandParams& p, int C_rt, int lane, int half, int nr, int img,
                                        uint32_t r0, const LgEmit& e, float lthr) {
  const int C = CT > 0 ? CT : C_rt;
  const float ninf = -__int_as_float(0x7f800000);
  const int q = lane & 3;
  const int lrow = 8 * half + (lane >> 2);
  const bool valid = lrow < nr;
  const float* row = tile + (valid ? lrow : nr - 1) * C;  // lanes past the last row redo it and flag nothing
  const float kLog2e = 1.4426950408889634f;
  float mfg, mx, sum;
  if (C <= 4 * kLgRegCols && p.col0 <= 1) {
    // The lane's quarter of the row is read from shared memory ONCE into registers.
    float v[kLgRegCols];
#pragma unroll
    for (int k = 0; k < kLgRegCols; ++k) {
      const int c = q + 4 * k;
      v[k] = c < C ? row[c] : ninf;
    }
    float m0 = ninf, m1 = ninf;
#pragma unroll
    for (int k = 1; k < kLgRegCols; k += 2) {
      m0 = fmaxf(m0, v[k]);
      if (k + 1 < kLgRegCols) m1 = fmaxf(m1, v[k + 1]);
    }
    const float rest = fmaxf(m0, m1);
    mx = fmaxf(rest, v[0]);
    // foreground maximum: everything but column 0 when col0 == 1 (element k = 0 of lane q = 0)
    float mf = (p.col0 == 1 && q == 0) ? rest : mx;
    mf = fmaxf(mf, __shfl_xor_sync(0xffffffffu, mf, 1));
    mfg = fmaxf(mf, __shfl_xor_sync(0xffffffffu, mf, 2));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    // The softmax denominator is at least 1 (the row maximum contributes exp(0)), so probability > threshold needs
    // x - max > log(threshold): a test on the two maxima alone.  Rows of a confident background fail it, and their
    // denominator is never formed.  The few rows that pass get theirs from the whole warp, three columns per lane, whose
    // terms double as the candidate test (term > threshold * sum, with a margin far above ex2.approx's error); only real
    // candidates pay for the exact expf and the division.  A half tile where many rows pass (an undecided head) runs
    // the register-resident pass below instead.
    const unsigned pass = __ballot_sync(0xffffffffu, valid && q == 0 && mfg - mx > lthr);
    if (__popc(pass) <= kLgWideRows) {
      unsigned todo = pass;
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const float rmx = __shfl_sync(0xffffffffu, mx, src);
        const int rrow = 8 * half + (src >> 2);
        const float* frow = tile + rrow * C;
        const float rn2 = -rmx * kLog2e;
        float ev[3];
        float part = 0.0f;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const int k = lane + 32 * j;
          ev[j] = k < C ? exp2f_approx(fmaf(frow[k], kLog2e, rn2)) : 0.0f;
          part += ev[j];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        const float rsum = part;
        const float cut = p.threshold * rsum * (1.0f - 2e-4f);
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const int k = lane + 32 * j;
          if (k < p.col0 || k >= C || !(ev[j] > cut)) continue;
          const float pr = expf(frow[k] - rmx) / rsum;
          if (!(pr > p.threshold)) continue;
          lg_emit_key(e, img, make_cand_key((uint32_t)(k - p.col0), pr, r0 + (uint32_t)rrow));
        }
      }
      return;
    }
    const float nm2 = -mx * kLog2e;
    float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
    for (int k = 0; k < kLgRegCols; k += 2) {
      s0 += exp2f_approx(fmaf(v[k], kLog2e, nm2));  // padding is -inf: ex2 gives 0
      if (k + 1 < kLgRegCols) s1 += exp2f_approx(fmaf(v[k + 1], kLog2e, nm2));
    }
    sum = s0 + s1;
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);
    // Candidate test on the register-resident columns, every row's four lanes for themselves: x - max > lim is necessary
    // (with margin) for expf(x - max) / sum > threshold.  Inside an object's run of anchors all eight rows of the warp
    // are candidate rows; scanning them one after the other with the whole warp cost ~100 instructions per row.
    const float lim = p.threshold > 0.0f ? logf(p.threshold * sum) - 1e-4f : ninf;
    unsigned hits = 0;
#pragma unroll
    for (int k = 0; k < kLgRegCols; ++k)
      if (v[k] - mx > lim) hits |= 1u << k;  // padding columns are -inf
    if (!valid) hits = 0;
    if (q < p.col0) hits &= ~1u;  // col0 <= 1 here: column 0 is element 0 of lane q = 0
    while (hits) {
      const int k = __ffs(hits) - 1;
      hits &= hits - 1;
      const int c = q + 4 * k;
      const float pr = expf(row[c] - mx) / sum;
      if (pr > p.threshold) lg_emit_key(e, img, make_cand_key((uint32_t)(c - p.col0), pr, r0 + (uint32_t)lrow));
    }
    return;
  } else {  // wider rows: two passes over shared memory
    const int cf = p.col0 + ((q - p.col0) & 3);  // first foreground column of this lane's residue
    float m0 = ninf, m1 = ninf;
    int c = cf;
    for (; c + 4 < C; c += 8) {
      m0 = fmaxf(m0, row[c]);
      m1 = fmaxf(m1, row[c + 4]);
    }
    if (c < C) m0 = fmaxf(m0, row[c]);
    mfg = fmaxf(m0, m1);
    mx = mfg;
    for (int k = q; k < p.col0; k += 4) mx = fmaxf(mx, row[k]);
    mfg = fmaxf(mfg, __shfl_xor_sync(0xffffffffu, mfg, 1));
    mfg = fmaxf(mfg, __shfl_xor_sync(0xffffffffu, mfg, 2));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    // sum of exp(x - max) as ex2.approx(x * log2(e) - max * log2(e)), one FFMA + one MUFU per element.  Terms that matter
    // to the sum have |x - max| of a few units, where this is within ~3e-7 of expf; the exact expf is kept for the
    // candidates' own probabilities.
    const float nm2 = -mx * kLog2e;
    float s0 = 0.0f, s1 = 0.0f;
    c = q;
    for (; c + 4 < C; c += 8) {
      s0 += exp2f_approx(fmaf(row[c], kLog2e, nm2));
      s1 += exp2f_approx(fmaf(row[c + 4], kLog2e, nm2));
    }
    if (c < C) s0 += exp2f_approx(fmaf(row[c], kLog2e, nm2));
    sum = s0 + s1;
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);
  }
  // x - max > lim is necessary (with margin) for expf(x - max) / sum > threshold; most rows have no such column.  The
  // flagged rows are scanned by the whole warp, 32 columns at a time.
  const float lim = p.threshold > 0.0f ? logf(p.threshold * sum) - 1e-4f : ninf;
  unsigned flagged = __ballot_sync(0xffffffffu, valid && q == 0 && mfg - mx > lim);
  while (flagged) {
    const int src = __ffs(flagged) - 1;
    flagged &= flagged - 1;
    const float rmx = __shfl_sync(0xffffffffu, mx, src);
    const float rsum = __shfl_sync(0xffffffffu, sum, src);
    const float rlim = __shfl_sync(0xffffffffu, lim, src);
    const int rrow = 8 * half + (src >> 2);
    const float* frow = tile + rrow * C;
    for (int k = p.col0 + lane; k < C; k += 32) {
      const float d = frow[k] - rmx;
      if (!(d > rlim)) continue;
      const float pr = expf(d) / rsum;
      if (!(pr > p.threshold)) continue;
      lg_emit_key(e, img, make_cand_key((uint32_t)(k - p.col0), pr, r0 + (uint32_t)rrow));
    }
  }
}

template <int CT>
__global__ void __launch_bounds__(kLgThreads, 4)
candidates_logits_kernel(CandParams p, int32_t* __restrict__ cand_count, uint64_t* __restrict__ cand_keys, int64_t cap,
                         int64_t key_stride, int tiles_per_img, int64_t total_floats) {
  asm volatile("griddepcontrol.launch_dependents;");  // fast_nms_kernel may be queued behind this grid (fused path)
  extern __shared__ __align__(128) unsigned char lg_smem[];  // [pairs][kLgBufs] tile buffers
  __shared__ __align__(8) uint64_t full[kLgPairs][kLgBufs];   // tile landed (bulk copy's byte count)
  __shared__ __align__(8) uint64_t empty[kLgPairs][kLgBufs];  // both warps of the pair are done reading it
  __shared__ LgWarpStage stages[kLgWarps];
  const int C = CT > 0 ? CT : p.ncols;
  const int tile_bytes = ((kLgRows * C * 4 + 15) & ~15) + 16;  // widened to 16-byte boundaries on both sides
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, pair = warp >> 1, half = warp & 1;
  const int total_tiles = p.batch * tiles_per_img;
  const float* base = p.scores;
  if (tid < kLgPairs * kLgBufs) {
    mbar_init(&full[tid / kLgBufs][tid % kLgBufs], 1);
    mbar_init(&empty[tid / kLgBufs][tid % kLgBufs], 2);
  }
  if (tid < kLgWarps) stages[tid].n = 0;
  __syncthreads();

  unsigned char* wbuf = lg_smem + (size_t)pair * kLgBufs * tile_bytes;
  LgEmit em;
  em.sg = &stages[warp];
  em.cand_count = cand_count;
  em.cand_keys = cand_keys;
  em.cap = cap;
  em.key_stride = key_stride;
  const bool producer = half == 0 && lane == 0;
  // the logits are read once: evict-first in the L2, like K1's ld.global.cs (in the step this kernel starts while ~100 MB
  // of the previous call's mask bytes are still dirty in the L2; with the default policy its 400 MB of reads push them
  // out and wait for the write-back)
  const uint64_t stream_policy = l2_evict_first_policy();

  auto issue = [&](int b, int t, int stage) {  // producer lane only; tile t of image b
    const int r0 = t * kLgRows, nr = min(kLgRows, p.rows - r0);
    const int64_t f0 = ((int64_t)b * p.rows + r0) * C, f1 = f0 + (int64_t)nr * C;
    const int64_t a0 = f0 & ~(int64_t)3;                                          // floats, 16-byte boundary below
    const int64_t a1 = min((f1 + 3) & ~(int64_t)3, total_floats & ~(int64_t)3);   // never past the tensor
    const uint32_t bytes = (uint32_t)((a1 - a0) * 4);
    mbar_expect_tx(&full[pair][stage], bytes);
    bulk_g2s_stream(wbuf + (size_t)stage * tile_bytes, base + a0, bytes, &full[pair][stage], stream_policy);
  };

  // every pair of warps owns a run of consecutive tiles
  const int gp = blockIdx.x * kLgPairs + pair, np = gridDim.x * kLgPairs;
  const int per_pair = (total_tiles + np - 1) / np;
  const int first = gp * per_pair, last = min(first + per_pair, total_tiles);
  int b = first / tiles_per_img, t = first - b * tiles_per_img;  // (image, tile in image), advanced without divisions
  int cur_img = -1;
  const float lthr = p.threshold > 0.0f ? logf(p.threshold) - 1e-4f : -__int_as_float(0x7f800000);
  // kLgBufs buffers per pair: tiles it + 1 .. it + kLgBufs - 1 are in flight while tile it is reduced
  int pb = b, pt = t;  // tile the producer issues next
  auto advance = [&](int& ib, int& itl) { if (++itl == tiles_per_img) { itl = 0; ++ib; } };
  if (producer) {
#pragma unroll
    for (int j = 0; j < kLgBufs - 1; ++j)
      if (first + j < last) { issue(pb, pt, j); advance(pb, pt); }
  }
  int it = 0, stage = 0, use = 0;  // stage = it % kLgBufs, use = it / kLgBufs
  for (int tile = first; tile < last; ++tile, ++it) {
    if (producer && tile + kLgBufs - 1 < last) {
      // the buffer of tile it + kLgBufs - 1 held tile it - 1: this warp left it at its last __syncwarp, the partner says so on `empty`
      const int ps = stage == 0 ? kLgBufs - 1 : stage - 1;
      if (it >= 1) mbar_wait(&empty[pair][ps], (stage == 0 ? use - 1 : use) & 1);
      issue(pb, pt, ps);
      advance(pb, pt);
    }
    if (b != cur_img) {
      if (cur_img >= 0) lg_flush(em, cur_img, lane);
      cur_img = b;
    }
    const int r0 = t * kLgRows, nr = min(kLgRows, p.rows - r0);
    const int64_t f0 = ((int64_t)b * p.rows + r0) * C;
    const int64_t a0 = f0 & ~(int64_t)3;
    const int64_t a1 = min((f0 + (int64_t)nr * C + 3) & ~(int64_t)3, total_floats & ~(int64_t)3);
    const int in_smem = (int)(a1 - f0);  // floats of this tile the bulk copy delivered (all but < 4 at the tensor's end)
    mbar_wait(&full[pair][stage], use & 1);
    if (nr * C <= in_smem)
      lg_tile<CT>(reinterpret_cast<const float*>(wbuf + (size_t)stage * tile_bytes) + (f0 - a0), p, C, lane, half, nr, b, (uint32_t)r0, em, lthr);
    else
      lg_tile<CT>(base + f0, p, C, lane, half, nr, b, (uint32_t)r0, em, lthr);
    // the generic-proxy reads of this stage must be ordered before the async-proxy write that refills it (WAR across proxies)
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[pair][stage]);
    if (em.sg->n > kLgStage - 32) lg_flush(em, cur_img, lane);  // keep room so that the direct path stays rare
    advance(b, t);
    if (++stage == kLgBufs) { stage = 0; ++use; }
  }
  if (cur_img >= 0) lg_flush(em, cur_img, lane);
}

int configure_candidates(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(candidates_logits_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  ISB_CUDA(h, cudaFuncSetAttribute(candidates_logits_kernel<81>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  ISB_CUDA(h, cudaFuncSetAttribute(candidates_logits_kernel<21>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  return ISB_OK;
}

int launch_candidates(isb_handle* h, const CandParams& p, cudaStream_t s) {
  if ((int64_t)p.rows * p.ncols >= (1ll << 31)) ISB_FAIL(h, ISB_ERR_LIMIT, "score tensor of one image exceeds 2^31 elements");
  ISB_CUDA(h, cudaMemsetAsync(h->cand_count + h->ws_b0, 0, sizeof(int32_t) * p.batch, s));
  if (p.logits) {
    const int tile_bytes = ((kLgRows * p.ncols * 4 + 15) & ~15) + 16;
    const size_t smem = (size_t)kLgPairs * kLgBufs * tile_bytes;
    if (smem > 100 * 1024)
      ISB_FAIL(h, ISB_ERR_LIMIT, "detect from logits supports at most %d classes", (100 * 1024 / (kLgBufs * kLgPairs) - 32) / (kLgRows * 4));
    if (!p.use_threshold) ISB_FAIL(h, ISB_ERR_BAD_ARG, "detect from logits needs a threshold");
    const int tiles_per_img = (p.rows + kLgRows - 1) / kLgRows;
    int64_t grid = (int64_t)p.batch * tiles_per_img;
    const int64_t per_sm = smem <= 50 * 1024 ? 4 : (smem <= 72 * 1024 ? 3 : 2);  // resident CTAs: 227 KB of shared memory
    if (grid > per_sm * h->num_sms) grid = per_sm * h->num_sms;
    auto kern = p.ncols == 81 ? candidates_logits_kernel<81> : (p.ncols == 21 ? candidates_logits_kernel<21> : candidates_logits_kernel<0>);
    kern<<<(unsigned)grid, kLgThreads, smem, s>>>(
        p, h->cand_count + h->ws_b0, h->cand_keys + (size_t)h->ws_b0 * h->cand_cap_pow2, h->lim.cand_capacity, h->cand_cap_pow2,
        tiles_per_img, (int64_t)p.batch * p.rows * p.ncols);
    ISB_LAUNCH_CHECK(h, "candidates_logits_kernel");
    return ISB_OK;
  }
  const int64_t nvec = ((int64_t)p.rows * p.ncols) / 4;
  int ctas = (int)((nvec + (int64_t)kCandThreads * kCandUnroll - 1) / ((int64_t)kCandThreads * kCandUnroll));
  // whole waves of 8 resident CTAs per SM; four waves measured best on B200 (76 us / 64 images vs 83 us for one)
  int per_img = (h->num_sms * 8 * 4) / p.batch;
  if (per_img < 1) per_img = 1;
  if (ctas > per_img) ctas = per_img;
  if (ctas < 1) ctas = 1;
  dim3 grid(ctas, p.batch);
  candidates_kernel<<<grid, kCandThreads, 0, s>>>(p, h->cand_count + h->ws_b0, h->cand_keys + (size_t)h->ws_b0 * h->cand_cap_pow2,
                                                  h->lim.cand_capacity, h->cand_cap_pow2);
  ISB_LAUNCH_CHECK(h, "candidates_kernel");
  return ISB_OK;
}

}  // namespace isb
