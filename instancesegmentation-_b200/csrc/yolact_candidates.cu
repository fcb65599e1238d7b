// K1 -- score filter + candidate emission.
//
// Replaces utils_bbox.py:193-204 (drop background column, row max over classes, > confidence,
// boolean gather) and the "transpose + per-class sort input" of :55-61.  Instead of gathering
// n x (C-1) score rows, every (row, class) entry whose score passes the threshold becomes one
// 64-bit key (class | ~score | row); see DESIGN.md "Exact shortcut" for why entries <= confidence
// can never change the result.
//
// HBM-bound: the score tensor [B, A, C] is read exactly once (6.2 MB / image at COCO 550).
// A CTA owns ROWS_PER_CTA consecutive rows = one contiguous byte range, pulled into shared memory
// by a single 1-D bulk-TMA copy (UBLKCP) signalled on an mbarrier; rows are then scanned from
// shared memory with an odd word stride (C = 81, 21, ...), which is bank-conflict free.
#include "isb_common.cuh"

namespace isb {

constexpr int kCandThreads = 128;  // one row per thread

__global__ void __launch_bounds__(kCandThreads)
candidates_kernel(CandParams p, int32_t* __restrict__ cand_count, uint64_t* __restrict__ cand_keys, int64_t cap,
                  int64_t key_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ int warp_tot[kCandThreads / 32];
  __shared__ int cta_base;

  const int b = blockIdx.y;
  const int row0 = blockIdx.x * kCandThreads;
  const int nrows = min(kCandThreads, p.rows - row0);
  const int tid = threadIdx.x;

  // contiguous byte range of this tile, widened to 16-byte boundaries
  const char* g0 = reinterpret_cast<const char*>(p.scores + ((size_t)b * p.rows + row0) * p.ncols);
  const uintptr_t ga = reinterpret_cast<uintptr_t>(g0);
  const uint32_t head = (uint32_t)(ga & 15);
  const uint32_t bytes = (head + (uint32_t)nrows * p.ncols * 4u + 15u) & ~15u;

  if (tid == 0) mbar_init(&bar, 1);
  __syncthreads();
  if (tid == 0) {
    mbar_expect_tx(&bar, bytes);
    bulk_g2s(smem_raw, g0 - head, bytes, &bar);
  }
  mbar_wait(&bar, 0);

  const float* tile = reinterpret_cast<const float*>(smem_raw + head);
  const float* row = tile + (size_t)tid * p.ncols;

  // pass 1: count this row's candidates
  int cnt = 0;
  if (tid < nrows) {
    if (p.use_threshold) {
      for (int c = p.col0; c < p.ncols; ++c) cnt += (row[c] > p.threshold) ? 1 : 0;
    } else {
      cnt = p.ncols - p.col0;
    }
  }
  // block exclusive scan
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((tid & 31) >= o) incl += v;
  }
  if ((tid & 31) == 31) warp_tot[tid >> 5] = incl;
  __syncthreads();
  int warp_off = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kCandThreads / 32; ++w) {
    int t = warp_tot[w];
    if (w < (tid >> 5)) warp_off += t;
    total += t;
  }
  if (total == 0) return;
  if (tid == 0) cta_base = atomicAdd(&cand_count[b], total);
  __syncthreads();
  int64_t pos = (int64_t)cta_base + warp_off + incl - cnt;

  // pass 2: emit keys (positions past the capacity are dropped; the NMS kernel reports overflow)
  if (tid < nrows && cnt > 0) {
    uint64_t* dst = cand_keys + (size_t)b * key_stride;
    for (int c = p.col0; c < p.ncols; ++c) {
      float s = row[c];
      if (!p.use_threshold || s > p.threshold) {
        if (pos < cap) dst[pos] = make_cand_key((uint32_t)(c - p.col0), s, (uint32_t)(row0 + tid));
        ++pos;
      }
    }
  }
}

static constexpr size_t kCandSmemMax = 200 * 1024;

int configure_candidates(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(candidates_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCandSmemMax));
  return ISB_OK;
}

int launch_candidates(isb_handle* h, const CandParams& p, cudaStream_t s) {
  // tile bytes + 16 of alignment head + 16 of round-up slack
  const size_t smem = (size_t)p.ncols * 4 * kCandThreads + 32;
  if (smem > kCandSmemMax) ISB_FAIL(h, ISB_ERR_LIMIT, "score rows of %d columns do not fit the candidate tile", p.ncols);
  ISB_CUDA(h, cudaMemsetAsync(h->cand_count + h->ws_b0, 0, sizeof(int32_t) * p.batch, s));
  dim3 grid((p.rows + kCandThreads - 1) / kCandThreads, p.batch);
  candidates_kernel<<<grid, kCandThreads, smem, s>>>(p, h->cand_count + h->ws_b0, h->cand_keys + (size_t)h->ws_b0 * h->cand_cap_pow2,
                                                     h->lim.cand_capacity,
                                                     h->cand_cap_pow2);
  ISB_LAUNCH_CHECK(h, "candidates_kernel");
  return ISB_OK;
}

}  // namespace isb
