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
