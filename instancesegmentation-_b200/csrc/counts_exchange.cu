// The path's one multi-GPU exchange -- every rank learns every image's detection count -- fused into the
// kernel that produces the counts, over NVLink peer memory.
//
// Round 1 posted an NCCL all-gather of int32[B_local] after every step: a separate kernel that needs an SM while the
// persistent mask kernel holds all of them, plus a cross-rank stream dependency (0.944 scaling efficiency at 8 GPUs).
// Here each rank owns a ring [depth][world * B_local] of 64-bit words in its own HBM, exported once through CUDA IPC and
// mapped by every peer.  The tail of fast_nms_kernel (one CTA per image) stores  epoch << 32 | count  for its image into
// the current ring slot of EVERY rank -- `world` 8-byte peer stores per image, no extra launch, no collective.  A word
// carries its own epoch, so no flag or ordering between words is needed; a reader waits (bounded) until all words of a
// slot show the epoch it wants.
#include "isb_common.cuh"

namespace isb {

constexpr int kCountsWaitThreads = 1024;

// status: 0 ok, 1 timed out (a peer never published this epoch), 2 a newer epoch already overwrote the slot
__global__ void __launch_bounds__(kCountsWaitThreads)
counts_wait_kernel(const unsigned long long* __restrict__ slot, int n, unsigned epoch, int32_t* __restrict__ counts,
                   int64_t* __restrict__ offsets, int32_t* __restrict__ status, long long timeout_ns) {
  __shared__ int s_scan[32];
  __shared__ long long s_run;
  __shared__ int s_bad;
  const int tid = threadIdx.x;
  if (tid == 0) { s_run = 0; s_bad = 0; }
  __syncthreads();
  long long t0;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  for (int base = 0; base < n; base += kCountsWaitThreads) {
    const int i = base + tid;
    int c = 0;
    if (i < n) {
      for (;;) {
        unsigned long long w;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(slot + i) : "memory");
        const unsigned e = (unsigned)(w >> 32);
        if (e == epoch) { c = (int)(unsigned)(w & 0xffffffffu); break; }
        if ((int)(e - epoch) > 0) { atomicMax(&s_bad, 2); break; }
        long long t1;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
        if (t1 - t0 > timeout_ns) { atomicMax(&s_bad, 1); break; }
        __nanosleep(200);
      }
      counts[i] = c;
    }
    const int v = c > 0 ? c : 0;  // -1 (candidate overflow) counts as no detections in the offsets
    int total;
    const int excl = cta_exclusive_scan(v, s_scan, &total);
    if (offsets && i < n) offsets[i] = s_run + excl;
    __syncthreads();
    if (tid == 0) s_run += total;
    __syncthreads();
  }
  if (tid == 0) *status = s_bad;
}

int launch_counts_wait(isb_handle* h, const unsigned long long* slot, int n, unsigned epoch, int32_t* counts, int64_t* offsets,
                       int32_t* status, long long timeout_ns, cudaStream_t s) {
  counts_wait_kernel<<<1, kCountsWaitThreads, 0, s>>>(slot, n, epoch, counts, offsets, status, timeout_ns);
  ISB_LAUNCH_CHECK(h, "counts_wait_kernel");
  return ISB_OK;
}

}  // namespace isb
