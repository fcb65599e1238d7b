// Host threads that expand the bit-packed mask transport into the caller's bool [H,W,m] bytes (see host_expand.h).
// Pure host C++ (g++): AVX-512BW / AVX2 / scalar kernels chosen once at run time, 64-byte non-temporal stores, a
// small worker pool fed with cache-line-aligned slices.
#include "host_expand.h"

#include <immintrin.h>
#include <sched.h>
#include <stdlib.h>
#include <string.h>

#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

namespace isb {

namespace {

constexpr size_t kSliceBytes = 256 * 1024;  // output bytes per work item (a multiple of 64)

// 64 bits starting at bit `bit0` of the little-endian bit stream `src`
inline uint64_t bits64(const uint64_t* src, size_t bit0) {
  const size_t w = bit0 >> 6;
  const unsigned k = (unsigned)(bit0 & 63);
  return k ? (src[w] >> k) | (src[w + 1] << (64 - k)) : src[w];
}

inline void expand_bytes_scalar(const uint64_t* src, size_t bit0, uint8_t* dst, size_t n) {
  for (size_t i = 0; i < n; ++i) {
    const size_t b = bit0 + i;
    dst[i] = (uint8_t)((src[b >> 6] >> (b & 63)) & 1u);
  }
}

// dst 64-byte aligned, n a multiple of 64
void expand_lines_scalar(const uint64_t* src, size_t bit0, uint8_t* dst, size_t n) {
  for (size_t i = 0; i < n; i += 64) {
    const uint64_t w = bits64(src, bit0 + i);
    uint64_t* o = reinterpret_cast<uint64_t*>(dst + i);
    for (int b = 0; b < 8; ++b) {
      const uint64_t x = ((w >> (8 * b)) & 0xffu) * 0x0101010101010101ull & 0x8040201008040201ull;  // bit t of the byte -> byte t
      o[b] = ((x + 0x7f7f7f7f7f7f7f7full) >> 7) & 0x0101010101010101ull;                              // non-zero byte -> 1
    }
  }
}

__attribute__((target("avx2"))) void expand_lines_avx2(const uint64_t* src, size_t bit0, uint8_t* dst, size_t n) {
  const __m256i bitsel = _mm256_set1_epi64x((long long)0x8040201008040201ull);
  const __m256i shuf = _mm256_setr_epi8(0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 3, 3, 3, 3, 3, 3, 3, 3);
  const __m256i one = _mm256_set1_epi8(1);
  for (size_t i = 0; i < n; i += 64) {
    const uint64_t w = bits64(src, bit0 + i);
    for (int h = 0; h < 2; ++h) {
      __m256i v = _mm256_shuffle_epi8(_mm256_set1_epi32((int)(uint32_t)(w >> (32 * h))), shuf);
      v = _mm256_and_si256(_mm256_cmpeq_epi8(_mm256_and_si256(v, bitsel), bitsel), one);
      _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 32 * h), v);
    }
  }
  _mm_sfence();
}

__attribute__((target("avx512f,avx512bw"))) void expand_lines_avx512(const uint64_t* src, size_t bit0, uint8_t* dst, size_t n) {
  const __m512i one = _mm512_set1_epi8(1);
  const unsigned k = (unsigned)(bit0 & 63);
  const uint64_t* s = src + (bit0 >> 6);
  if (k == 0) {
    for (size_t i = 0; i < n; i += 64)
      _mm512_stream_si512(reinterpret_cast<__m512i*>(dst + i), _mm512_maskz_mov_epi8((__mmask64)s[i >> 6], one));
  } else {
    uint64_t lo = s[0];
    for (size_t i = 0; i < n; i += 64) {
      const uint64_t hi = s[(i >> 6) + 1];
      _mm512_stream_si512(reinterpret_cast<__m512i*>(dst + i), _mm512_maskz_mov_epi8((__mmask64)((lo >> k) | (hi << (64 - k))), one));
      lo = hi;
    }
  }
  _mm_sfence();
}

// experiment: regular (write-allocate) 64-byte stores instead of non-temporal ones
__attribute__((target("avx512f,avx512bw"))) void expand_lines_avx512_cached(const uint64_t* src, size_t bit0, uint8_t* dst, size_t n) {
  const __m512i one = _mm512_set1_epi8(1);
  for (size_t i = 0; i < n; i += 64)
    _mm512_store_si512(reinterpret_cast<__m512i*>(dst + i), _mm512_maskz_mov_epi8((__mmask64)bits64(src, bit0 + i), one));
}
void expand_lines_dry(const uint64_t*, size_t, uint8_t*, size_t) {}  // experiment: no stores at all

using LinesFn = void (*)(const uint64_t*, size_t, uint8_t*, size_t);

struct Isa {
  LinesFn fn;
  const char* name;
};

Isa pick_isa() {
  const char* force = getenv("ISB_HOST_ISA");
  __builtin_cpu_init();
  const bool a512 = __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512f");
  const bool a2 = __builtin_cpu_supports("avx2");
  if (force && !strcmp(force, "scalar")) return {expand_lines_scalar, "scalar"};
  if (force && !strcmp(force, "dry")) return {expand_lines_dry, "dry (experiment: nothing written)"};
  if (force && !strcmp(force, "cached") && a512) return {expand_lines_avx512_cached, "avx512bw, cached stores (experiment)"};
  if (force && !strcmp(force, "avx2") && a2) return {expand_lines_avx2, "avx2"};
  if (a512) return {expand_lines_avx512, "avx512bw"};
  if (a2) return {expand_lines_avx2, "avx2"};
  return {expand_lines_scalar, "scalar"};
}

const Isa& isa() {
  static const Isa i = pick_isa();
  return i;
}

// output byte range [a, b) of a job: unaligned head / tail bytes one by one, the cache-line-aligned interior as lines
void expand_range(const uint64_t* src, uint8_t* dst, size_t a, size_t b) {
  if (a >= b) return;
  size_t head = (size_t)((64 - (reinterpret_cast<uintptr_t>(dst + a) & 63)) & 63);
  if (head > b - a) head = b - a;
  expand_bytes_scalar(src, a, dst + a, head);
  const size_t body = (b - a - head) & ~(size_t)63;
  if (body) isa().fn(src, a + head, dst + a + head, body);
  expand_bytes_scalar(src, a + head + body, dst + a + head + body, b - a - head - body);
}

// slice s of a job: slices are cut at cache-line boundaries of the DESTINATION so that no two threads share a line
void slice_bounds(const ExpandJob* j, int s, size_t* a, size_t* b) {
  const size_t lead = (size_t)((64 - (reinterpret_cast<uintptr_t>(j->dst) & 63)) & 63);  // bytes before the first aligned line
  *a = s == 0 ? 0 : lead + (size_t)s * kSliceBytes;
  *b = lead + (size_t)(s + 1) * kSliceBytes;
  if (*b > j->out_bytes || s == j->nslices - 1) *b = j->out_bytes;
  if (*a > j->out_bytes) *a = j->out_bytes;
}

}  // namespace

struct HostPool {
  std::vector<std::thread> workers;
  std::mutex mu;
  std::condition_variable cv;
  std::deque<ExpandJob*> jobs;
  bool stop = false;
};

static void worker_main(HostPool* p) {
  for (;;) {
    ExpandJob* job = nullptr;
    int s = -1;
    {
      std::unique_lock<std::mutex> lk(p->mu);
      for (;;) {
        if (p->stop) return;
        if (!p->jobs.empty()) {
          job = p->jobs.front();
          s = job->next.fetch_add(1, std::memory_order_relaxed);
          if (s < job->nslices) break;
          p->jobs.pop_front();  // every slice has an owner: later workers move on to the next job
          continue;
        }
        p->cv.wait(lk);
      }
    }
    size_t a, b;
    slice_bounds(job, s, &a, &b);
    expand_range(job->src, job->dst, a, b);
    job->finished.fetch_add(1, std::memory_order_release);
  }
}

HostPool* host_pool_create(int threads) {
  HostPool* p = new HostPool();
  if (threads < 1) threads = 1;
  for (int t = 0; t < threads; ++t) p->workers.emplace_back(worker_main, p);
  return p;
}

void host_pool_destroy(HostPool* p) {
  if (!p) return;
  {
    std::lock_guard<std::mutex> lk(p->mu);
    p->stop = true;
  }
  p->cv.notify_all();
  for (auto& t : p->workers) t.join();
  delete p;
}

int host_pool_threads(const HostPool* p) { return p ? (int)p->workers.size() : 0; }
const char* host_pool_isa(const HostPool*) { return isa().name; }

void host_pool_submit(HostPool* p, ExpandJob* job) {
  const size_t lead = (size_t)((64 - (reinterpret_cast<uintptr_t>(job->dst) & 63)) & 63);
  const size_t rest = job->out_bytes > lead ? job->out_bytes - lead : 0;
  {
    std::lock_guard<std::mutex> lk(p->mu);
    // a finished job stays queued until a worker walks past it: drop such a stale entry before the object is reused
    for (auto it = p->jobs.begin(); it != p->jobs.end();) it = (*it == job) ? p->jobs.erase(it) : it + 1;
    job->nslices = (int)((rest + kSliceBytes - 1) / kSliceBytes);
    if (job->nslices < 1) job->nslices = 1;
    job->next.store(0, std::memory_order_relaxed);
    job->finished.store(0, std::memory_order_relaxed);
    p->jobs.push_back(job);
  }
  p->cv.notify_all();
}

void host_expand_serial(const uint64_t* src, uint8_t* dst, size_t out_bytes) { expand_range(src, dst, 0, out_bytes); }

int host_default_threads() {
  if (const char* e = getenv("ISB_HOST_THREADS")) return atoi(e);
  cpu_set_t set;
  int n = 0;
  if (sched_getaffinity(0, sizeof(set), &set) == 0) n = CPU_COUNT(&set);
  if (n <= 0) n = (int)std::thread::hardware_concurrency();
  int ranks = 1;
  if (const char* e = getenv("LOCAL_WORLD_SIZE")) ranks = atoi(e) > 0 ? atoi(e) : 1;
  n = n / ranks - 1;
  // more workers than ~8 write faster than the host's memory system can absorb next to the upload's DMA reads: measured on
  // the 16-vCPU B200 box (profiles/r02_host_delivery.json): 22.8 ms per 64 images with 8 workers, 24.9 ms with 15
  if (n > 8) n = 8;
  return n < 1 ? 1 : n;
}

}  // namespace isb
