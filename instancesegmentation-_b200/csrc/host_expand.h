// Host side of the bit-packed mask transport of isb_yolact_detect_host (internal, not part of the ABI).
//
// The device writes bool [H,W,m] one BIT per element (mask_compose_kernel<kCompPacked>); an eighth of the bytes cross
// PCIe, and a pool of host threads expands them into the caller's byte-per-element buffer with non-temporal stores
// while the copy engine delivers other images densely -- the caller's contract (host bool [H,W,m]) is unchanged.
// This is a transport decode (bit i -> byte i), not a compute fallback: every mask bit is computed on the GPU.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <atomic>

namespace isb {

struct HostPool;

// One expansion in flight: `out_bytes` bits from `src` (8-byte aligned; readable up to 16 bytes past the last bit) become
// `out_bytes` bytes (0 / 1) at `dst` (any alignment).  Owned by the caller; reusable once done() is true.
struct ExpandJob {
  const uint64_t* src = nullptr;
  uint8_t* dst = nullptr;
  size_t out_bytes = 0;
  int nslices = 0;
  std::atomic<int> next{0};
  std::atomic<int> finished{0};
  bool done() const { return finished.load(std::memory_order_acquire) >= nslices; }
};

HostPool* host_pool_create(int threads);
void host_pool_destroy(HostPool* p);
int host_pool_threads(const HostPool* p);
const char* host_pool_isa(const HostPool* p);  // "avx512bw" | "avx2" | "scalar"
// Queue the job for the workers (returns at once).  The job must stay alive and untouched until done().
void host_pool_submit(HostPool* p, ExpandJob* job);
// Expand on the calling thread (tests, tiny inputs).
void host_expand_serial(const uint64_t* src, uint8_t* dst, size_t out_bytes);
// Default worker count: CPUs this process may run on (sched_getaffinity), divided by LOCAL_WORLD_SIZE when a launcher
// set it (one process per GPU shares the host), minus one for the thread that drives the copies, at most 8 (measured optimum);
// ISB_HOST_THREADS overrides.
int host_default_threads();

}  // namespace isb
