// Write-only HBM bandwidth of a B200 under different store strategies (context for the roofline of the
// write-bound kernels: YOLACT mask assembly, Mask R-CNN paste).  Stand-alone:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/write_bw tools/write_bw.cu && /tmp/write_bw
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

enum { PLAIN = 0, CS = 1, NOALLOC = 2, V8 = 3, EVICT_FIRST = 4, WT = 5 };

template <int MODE>
__device__ __forceinline__ void store16(uint4* p, uint64_t pol) {
  if (MODE == PLAIN) *p = make_uint4(0, 0, 0, 0);
  if (MODE == CS) asm volatile("st.global.cs.v4.u32 [%0], {%1,%1,%1,%1};" ::"l"(p), "r"(0) : "memory");
  if (MODE == WT) asm volatile("st.global.wt.v4.u32 [%0], {%1,%1,%1,%1};" ::"l"(p), "r"(0) : "memory");
  if (MODE == NOALLOC) asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%1,%1,%1};" ::"l"(p), "r"(0) : "memory");
  if (MODE == EVICT_FIRST)
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%1,%1,%1}, %2;" ::"l"(p), "r"(0), "l"(pol) : "memory");
}

// contiguous chunk per CTA iteration: CTA c writes [c*chunk, (c+1)*chunk) then strides by grid
template <int MODE>
__global__ void fill16(uint4* out, int64_t nvec, int chunk_vec) {
  uint64_t pol = 0;
  if (MODE == EVICT_FIRST) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  for (int64_t base = (int64_t)blockIdx.x * chunk_vec; base < nvec; base += (int64_t)gridDim.x * chunk_vec) {
    const int64_t end = base + chunk_vec < nvec ? base + chunk_vec : nvec;
    for (int64_t v = base + threadIdx.x; v < end; v += blockDim.x) store16<MODE>(out + v, pol);
  }
}

__global__ void fill32(uint4* out, int64_t nvec, int chunk_vec) {  // 256-bit stores
  for (int64_t base = (int64_t)blockIdx.x * chunk_vec; base < nvec; base += (int64_t)gridDim.x * chunk_vec) {
    const int64_t end = base + chunk_vec < nvec ? base + chunk_vec : nvec;
    for (int64_t v = base + 2 * threadIdx.x; v + 1 < end + 1 && v < end; v += 2 * blockDim.x)
      asm volatile("st.global.v8.u32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(out + v), "r"(0) : "memory");
  }
}

// TMA bulk stores from a zeroed shared buffer
__global__ void fill_bulk(uint8_t* out, int64_t nbytes, int chunk) {
  extern __shared__ __align__(128) uint8_t zbuf[];
  for (int i = threadIdx.x; i < chunk / 16; i += blockDim.x) reinterpret_cast<uint4*>(zbuf)[i] = make_uint4(0, 0, 0, 0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(zbuf);
    int inflight = 0;
    for (int64_t base = (int64_t)blockIdx.x * chunk; base < nbytes; base += (int64_t)gridDim.x * chunk) {
      const uint32_t n = (uint32_t)(base + chunk <= nbytes ? chunk : nbytes - base);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + base), "r"(s), "r"(n) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (++inflight >= 8) { asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory"); inflight = 4; }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

__global__ void copy16(const uint4* __restrict__ in, uint4* __restrict__ out, int64_t nvec) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) out[v] = in[v];
}
__global__ void read16(const uint4* __restrict__ in, int64_t nvec, uint32_t* sink) {
  uint32_t acc = 0;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    uint4 a = in[v];
    acc ^= a.x ^ a.y ^ a.z ^ a.w;
  }
  if (acc == 0x12345678u) *sink = acc;
}

template <typename F>
static float best_ms(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) f();
  float best = 1e9f;
  for (int i = 0; i < 10; ++i) {
    cudaEventRecord(e0);
    f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  const int64_t nbytes = 2200000000ll / 16 * 16;
  const int64_t nvec = nbytes / 16;
  uint8_t *a, *b;
  uint32_t* sink;
  CK(cudaMalloc(&a, nbytes));
  CK(cudaMalloc(&b, nbytes));
  CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(a, 1, nbytes));
  CK(cudaMemset(b, 1, nbytes));
  CK(cudaFuncSetAttribute(fill_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
  printf("{\n");
  float ms = best_ms([&] { cudaMemsetAsync(a, 0, nbytes); });
  printf(" \"cudaMemset\": %.1f,\n", nbytes / ms / 1e6);
  const int sms = 148;
  const int per_sm[] = {1, 2, 4, 8};
  const int threads[] = {256, 512, 1024};
  const int chunks[] = {4096, 65536};  // bytes per CTA-iteration
  for (int ps : per_sm)
    for (int th : threads)
      for (int ch : chunks) {
        if (ps * th > 2048) continue;
        const int grid = sms * ps, cv = ch / 16 * (th / 256);
#define RUN(name, MODE)                                                                                   \
  ms = best_ms([&] { fill16<MODE><<<grid, th>>>((uint4*)a, nvec, cv); });                                 \
  printf(" \"%s_ctas%d_t%d_chunk%d\": %.1f,\n", name, ps, th, cv * 16, nbytes / ms / 1e6);
        RUN("plain", PLAIN)
        RUN("cs", CS)
        RUN("noalloc", NOALLOC)
        RUN("evict_first", EVICT_FIRST)
        RUN("wt", WT)
        ms = best_ms([&] { fill32<<<grid, th>>>((uint4*)a, nvec, cv); });
        printf(" \"v8_ctas%d_t%d_chunk%d\": %.1f,\n", ps, th, cv * 16, nbytes / ms / 1e6);
      }
  for (int ps : {1, 2, 4})
    for (int ch : {8192, 32768, 65536}) {
      if (ps * ch > 200 * 1024) continue;
      ms = best_ms([&] { fill_bulk<<<sms * ps, 128, ch>>>(a, nbytes, ch); });
      printf(" \"bulk_s2g_ctas%d_chunk%d\": %.1f,\n", ps, ch, nbytes / ms / 1e6);
    }
  CK(cudaGetLastError());
  ms = best_ms([&] { copy16<<<sms * 8, 256>>>((const uint4*)a, (uint4*)b, nvec); });
  printf(" \"copy_rw\": %.1f,\n", 2.0 * nbytes / ms / 1e6);
  ms = best_ms([&] { cudaMemcpyAsync(b, a, nbytes, cudaMemcpyDeviceToDevice); });
  printf(" \"cudaMemcpy_rw\": %.1f,\n", 2.0 * nbytes / ms / 1e6);
  ms = best_ms([&] { read16<<<sms * 8, 256>>>((const uint4*)a, nvec, sink); });
  printf(" \"read_only\": %.1f\n}\n", nbytes / ms / 1e6);
  CK(cudaDeviceSynchronize());
  return 0;
}
