// Issue rate of dp4a (IDP.4A) against IMAD / PRMT / SHF on sm_100a: warp instructions per clock per SM sub-partition.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/idp_rate tools/idp_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int OP>
__global__ void rate_kernel(uint32_t* out, uint32_t seed, int iters, long long* cycles) {
  uint32_t a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed * (threadIdx.x + 1) + i;
  const uint32_t b = seed ^ 0x01020304u, c = seed | 7u;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (OP == 0) asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
        if (OP == 1) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
        if (OP == 2) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(0x3120u + r));
        if (OP == 3) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
        if (OP == 4) asm volatile("dp2a.lo.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
        if (OP == 5) asm volatile("bfe.u32 %0, %0, 8, 8;" : "+r"(a[i]));
        if (OP == 6) asm volatile("vmad.u32.u32.u32 %0, %1.b0, %2.b1, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
      }
  }
  const long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int OP>
void run(const char* name, uint32_t* out, long long* cyc) {
  const int iters = 4096, threads = 512;  // 16 warps per SM, 4 per sub-partition
  rate_kernel<OP><<<148, threads>>>(out, 12345u, iters, cyc);
  cudaDeviceSynchronize();
  rate_kernel<OP><<<148, threads>>>(out, 12345u, iters, cyc);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  const double instr_per_smsp = (double)iters * 32 * (threads / 32 / 4);
  printf("\"%s\": {\"cycles\": %lld, \"warp_instr_per_clk_per_smsp\": %.3f},\n", name, c, instr_per_smsp / (double)c);
}

int main() {
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 512 * 4);
  cudaMalloc(&cyc, 8);
  printf("{\n");
  run<0>("dp4a", out, cyc);
  run<1>("imad", out, cyc);
  run<2>("prmt", out, cyc);
  run<3>("shf", out, cyc);
  run<4>("dp2a", out, cyc);
  run<5>("bfe", out, cyc);
  run<6>("vmad", out, cyc);
  printf("\"err\": \"%s\"\n}\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
