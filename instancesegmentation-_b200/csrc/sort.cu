// Device-wide ascending sort of 64-bit keys (power-of-two length, padded with ~0).
//
// Used by the Mask R-CNN side to order boxes by score (boxes.py:35 -> torchvision nms sorts by score,
// stable) and objectness per level (rpn_function.py:455 topk).  Keys carry the original index in their
// low bits, so an unstable network still yields the stable order.
//
// Bitonic network, blocked for the memory hierarchy: every (size, stride) stage with stride < 4096
// runs inside shared memory on 4096-key chunks (32 KB), only the strides >= 4096 touch global memory
// one stage per launch.  For n <= 4096 the whole sort is a single launch.
#include "isb_common.cuh"

namespace isb {

constexpr int kSortChunk = 4096;
constexpr int kSortThreads = 1024;

__device__ __forceinline__ void cmpx(uint64_t& a, uint64_t& b, bool asc) {
  if ((a > b) == asc) { uint64_t t = a; a = b; b = t; }
}

// Sort each chunk completely (all sizes up to chunk length); direction taken from the GLOBAL index so
// that the chunks form the alternating runs the later merge stages expect.
__global__ void __launch_bounds__(kSortThreads) sort_chunks_kernel(uint64_t* __restrict__ keys, int n) {
  __shared__ uint64_t s[kSortChunk];
  const int len = min(kSortChunk, n);
  const int base = blockIdx.x * len;
  for (int i = threadIdx.x; i < len; i += kSortThreads) s[i] = keys[base + i];
  for (int size = 2; size <= len; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int t = threadIdx.x; t < (len >> 1); t += kSortThreads) {
        const int lo = 2 * t - (t & (stride - 1));
        const bool asc = ((base + lo) & size) == 0;
        cmpx(s[lo], s[lo + stride], asc);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < len; i += kSortThreads) keys[base + i] = s[i];
}

// One global compare-exchange stage (stride >= chunk).
__global__ void __launch_bounds__(256) sort_global_step_kernel(uint64_t* __restrict__ keys, int n, int size, int stride) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (n >> 1)) return;
  const int lo = 2 * t - (t & (stride - 1));
  const bool asc = (lo & size) == 0;
  uint64_t a = keys[lo], b = keys[lo + stride];
  if ((a > b) == asc) { keys[lo] = b; keys[lo + stride] = a; }
}

// Finish a merge level inside shared memory: strides chunk/2 .. 1 of the given size.
__global__ void __launch_bounds__(kSortThreads) sort_merge_chunks_kernel(uint64_t* __restrict__ keys, int size) {
  __shared__ uint64_t s[kSortChunk];
  const int base = blockIdx.x * kSortChunk;
  for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) s[i] = keys[base + i];
  for (int stride = kSortChunk >> 1; stride > 0; stride >>= 1) {
    __syncthreads();
    for (int t = threadIdx.x; t < (kSortChunk >> 1); t += kSortThreads) {
      const int lo = 2 * t - (t & (stride - 1));
      const bool asc = ((base + lo) & size) == 0;
      cmpx(s[lo], s[lo + stride], asc);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) keys[base + i] = s[i];
}

// ---- rank sort: for up to a few ten thousand UNIQUE keys ------------------------------------------------
// rank(i) = number of keys smaller than key i; out[rank(i)] = key i.  O(n^2) compares, but they are perfectly
// parallel, barrier-free and need one launch each for counting and scattering: 23 M compares for the 4819 RPN
// candidates take a few microseconds, where the bitonic network needs ~90 dependent stages.  grid (i-blocks, j-slices):
// a CTA counts, for its 256 keys, the smaller keys inside its slice of the array (staged through shared memory)
// and adds the partial count to rank[i].
constexpr int kRankThreads = 256;

__global__ void __launch_bounds__(kRankThreads) rank_count_kernel(const uint64_t* __restrict__ keys, int n_max,
                                                                  const int32_t* __restrict__ n_dev, int slice,
                                                                  int32_t* __restrict__ rank) {
  __shared__ uint64_t tile[kRankThreads];
  pdl_enter();
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  const int i = blockIdx.x * kRankThreads + threadIdx.x;
  if (blockIdx.x * kRankThreads >= n) return;  // CTA-uniform
  const int jb = blockIdx.y * slice, je = min(n, jb + slice);
  if (jb >= je) return;
  const uint64_t mine = i < n ? keys[i] : 0ull;
  int cnt = 0;
  for (int j0 = jb; j0 < je; j0 += kRankThreads) {
    __syncthreads();
    const int j = j0 + threadIdx.x;
    tile[threadIdx.x] = j < je ? keys[j] : ~0ull;  // pad: never smaller than a real key
    __syncthreads();
#pragma unroll 8
    for (int t = 0; t < kRankThreads; ++t) cnt += tile[t] < mine ? 1 : 0;
  }
  if (i < n && cnt) atomicAdd(&rank[i], cnt);
}

__global__ void rank_scatter_kernel(const uint64_t* __restrict__ keys, int n_max, const int32_t* __restrict__ n_dev,
                                    const int32_t* __restrict__ rank, uint64_t* __restrict__ out) {
  pdl_enter();
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[rank[i]] = keys[i];
}

// keys [n] unique -> out [n] ascending; rank: int32 scratch [n_max].  n is min(*n_dev, n_max) when n_dev != nullptr.
int launch_rank_sort_u64(isb_handle* h, const uint64_t* keys, int n_max, const int32_t* n_dev, int32_t* rank, uint64_t* out,
                         cudaStream_t s, bool rank_is_zero) {
  if (n_max <= 0) return ISB_OK;
  // rank_is_zero: the caller cleared rank[0 .. n_max) earlier on the stream and a kernel of the same call is the stream's last
  // operation, so the count kernel joins the caller's programmatic-dependent chain instead of following a memset node
  if (!rank_is_zero) ISB_CUDA(h, cudaMemsetAsync(rank, 0, (size_t)n_max * sizeof(int32_t), s));
  const int iblocks = (n_max + kRankThreads - 1) / kRankThreads;
  int slices = (4 * h->num_sms + iblocks - 1) / iblocks;  // enough CTAs for ~4 per SM
  if (slices > iblocks) slices = iblocks;
  if (slices < 1) slices = 1;
  int slice = (n_max + slices - 1) / slices;
  slice = (slice + kRankThreads - 1) / kRankThreads * kRankThreads;
  slices = (n_max + slice - 1) / slice;
  const cudaError_t lc = launch_chained(rank_count_kernel, dim3(iblocks, slices), dim3(kRankThreads), 0, s, rank_is_zero, keys, n_max,
                                        n_dev, slice, rank);
  if (lc != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "launch rank_count_kernel: %s", cudaGetErrorString(lc));
  h->launches++;
  const int sblocks = (n_max + 255) / 256 < 4 * h->num_sms ? (n_max + 255) / 256 : 4 * h->num_sms;
  const cudaError_t le = launch_chained(rank_scatter_kernel, dim3(sblocks), dim3(256), 0, s, true, keys, n_max, n_dev,
                                        (const int32_t*)rank, out);  // right behind rank_count_kernel
  if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "launch rank_scatter_kernel: %s", cudaGetErrorString(le));
  h->launches++;
  return ISB_OK;
}

int launch_sort_u64(isb_handle* h, uint64_t* keys, int n_pow2, cudaStream_t s) {
  if (n_pow2 <= 1) return ISB_OK;
  if (n_pow2 & (n_pow2 - 1)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "sort length %d is not a power of two", n_pow2);
  const int chunks = n_pow2 <= kSortChunk ? 1 : n_pow2 / kSortChunk;
  sort_chunks_kernel<<<chunks, kSortThreads, 0, s>>>(keys, n_pow2);
  ISB_LAUNCH_CHECK(h, "sort_chunks_kernel");
  for (int size = 2 * kSortChunk; size <= n_pow2; size <<= 1) {
    for (int stride = size >> 1; stride >= kSortChunk; stride >>= 1) {
      sort_global_step_kernel<<<(n_pow2 / 2 + 255) / 256, 256, 0, s>>>(keys, n_pow2, size, stride);
      ISB_LAUNCH_CHECK(h, "sort_global_step_kernel");
    }
    sort_merge_chunks_kernel<<<chunks, kSortThreads, 0, s>>>(keys, size);
    ISB_LAUNCH_CHECK(h, "sort_merge_chunks_kernel");
  }
  return ISB_OK;
}

}  // namespace isb
