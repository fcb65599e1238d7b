// Probe for the host-delivery design of isb_yolact_detect_host: how fast can T host threads expand a bit stream into a
// byte-per-bit stream in PINNED memory (non-temporal stores), alone and while the GPU's copy engine writes the same
// host memory at PCIe rate?   nvcc -O3 -o tools/bin/host_expand_bw tools/host_expand_bw.cu -Xcompiler -pthread,-mavx2
#include <cuda_runtime.h>
#include <immintrin.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

__attribute__((target("avx512f,avx512bw"))) static void expand512(const uint64_t* src, uint8_t* dst, size_t nwords) {
  const __m512i one = _mm512_set1_epi8(1);
  for (size_t i = 0; i < nwords; ++i) {
    const __m512i v = _mm512_maskz_mov_epi8((__mmask64)src[i], one);
    _mm512_stream_si512((__m512i*)(dst + i * 64), v);
  }
  _mm_sfence();
}
__attribute__((target("avx2"))) static void expand256(const uint64_t* src, uint8_t* dst, size_t nwords) {
  const __m256i bitsel = _mm256_set1_epi64x(0x8040201008040201ull);
  const __m256i shuf = _mm256_setr_epi8(0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 3, 3, 3, 3, 3, 3, 3, 3);
  const __m256i one = _mm256_set1_epi8(1);
  for (size_t i = 0; i < nwords; ++i) {
    const uint64_t w = src[i];
    for (int h = 0; h < 2; ++h) {
      const uint32_t x = (uint32_t)(w >> (32 * h));
      __m256i v = _mm256_shuffle_epi8(_mm256_set1_epi32((int)x), shuf);
      v = _mm256_and_si256(_mm256_cmpeq_epi8(_mm256_and_si256(v, bitsel), bitsel), one);
      _mm256_stream_si256((__m256i*)(dst + i * 64 + h * 32), v);
    }
  }
  _mm_sfence();
}

int main(int argc, char** argv) {
  const size_t out_bytes = (size_t)(argc > 1 ? atof(argv[1]) : 2.0) * (1ull << 30);
  cpu_set_t set;
  sched_getaffinity(0, sizeof(set), &set);
  const int ncpu = CPU_COUNT(&set);
  const bool has512 = __builtin_cpu_supports("avx512bw");
  printf("{\"affinity_cpus\": %d, \"hw_concurrency\": %u, \"avx512bw\": %s,\n", ncpu, std::thread::hardware_concurrency(), has512 ? "true" : "false");
  uint8_t *dst, *dma_dst;
  uint64_t* src;
  cudaHostAlloc(&dst, out_bytes, cudaHostAllocDefault);
  cudaHostAlloc(&dma_dst, 1ull << 30, cudaHostAllocDefault);
  cudaHostAlloc(&src, out_bytes / 8, cudaHostAllocDefault);
  for (size_t i = 0; i < out_bytes / 64; ++i) src[i] = (i % 31 == 0) ? 0x0000ff0000000101ull * i : 0;
  memset(dst, 1, out_bytes);
  void* dev;
  cudaMalloc(&dev, 1ull << 30);
  cudaMemset(dev, 1, 1ull << 30);
  cudaStream_t st;
  cudaStreamCreate(&st);
  const int tries[] = {1, 2, 4, 8, 12, 16, 24, 32, 48, 64};
  printf(" \"runs\": [\n");
  bool first = true;
  for (int with_dma = 0; with_dma < 2; ++with_dma)
    for (int ti = 0; ti < 10; ++ti) {
      const int T = tries[ti];
      if (T > 2 * ncpu) break;
      std::atomic<bool> stop{false};
      std::atomic<long> dma_bytes{0};
      std::thread dma;
      if (with_dma)
        dma = std::thread([&] {
          while (!stop.load()) {
            cudaMemcpyAsync(dma_dst, dev, 1ull << 30, cudaMemcpyDeviceToHost, st);
            cudaStreamSynchronize(st);
            dma_bytes += 1l << 30;
          }
        });
      double best = 1e30;
      const auto t_all0 = std::chrono::steady_clock::now();
      for (int rep = 0; rep < 3; ++rep) {
        const auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> th;
        const size_t words = out_bytes / 64, per = (words + T - 1) / T;
        for (int t = 0; t < T; ++t)
          th.emplace_back([=] {
            const size_t a = t * per, b = a + per < words ? a + per : words;
            if (a >= b) return;
            if (has512) expand512(src + a, dst + a * 64, b - a); else expand256(src + a, dst + a * 64, b - a);
          });
        for (auto& x : th) x.join();
        const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (s < best) best = s;
      }
      const double s_all = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_all0).count();
      stop = true;
      if (with_dma) dma.join();
      const double s_tot = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_all0).count();
      printf("%s  {\"threads\": %d, \"concurrent_d2h\": %s, \"expand_GBps\": %.1f, \"d2h_GBps_meanwhile\": %.1f}", first ? "" : ",\n", T,
             with_dma ? "true" : "false", out_bytes / best / 1e9, with_dma ? dma_bytes.load() / s_tot / 1e9 : 0.0);
      first = false;
      (void)s_all;
    }
  printf("\n ]}\n");
  // correctness spot check
  size_t bad = 0;
  for (size_t i = 0; i < 1 << 20; ++i) bad += dst[i] != ((src[i / 64] >> (i % 64)) & 1);
  fprintf(stderr, "mismatches in the first MiB: %zu\n", bad);
  return bad != 0;
}
