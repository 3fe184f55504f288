// ubench_int.cu -- integer-multiply throughput probes on sm_100a (what Philox4x32 is made of).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int ITER = 4096, ILP = 8;
template <int OP>
__global__ void __launch_bounds__(256) probe(uint32_t* out, uint32_t seed, int iters) {
  uint32_t a[ILP], b[ILP];
#pragma unroll
  for (int k = 0; k < ILP; ++k) { a[k] = threadIdx.x * 7u + k + seed; b[k] = a[k] ^ 0x9E3779B9u; }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < ILP; ++k) {
      if (OP == 0) { unsigned long long w; asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(a[k]), "r"(0xD2511F53u)); a[k] = (uint32_t)w ^ b[k]; b[k] = (uint32_t)(w >> 32); }   // IMAD.WIDE + LOP3
      if (OP == 1) { uint32_t h, l; asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(h) : "r"(a[k]), "r"(0xD2511F53u)); asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(l) : "r"(a[k]), "r"(0xD2511F53u)); a[k] = l ^ b[k]; b[k] = h; }   // IMAD.HI + IMAD + LOP3
      if (OP == 2) { uint32_t l; asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(l) : "r"(a[k]), "r"(0xD2511F53u)); a[k] = l ^ b[k]; }   // IMAD + LOP3
      if (OP == 3) { uint32_t h; asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(h) : "r"(a[k]), "r"(0xD2511F53u)); a[k] = h ^ b[k]; }   // IMAD.HI + LOP3
      if (OP == 4) { a[k] = (a[k] ^ b[k]) + 0x9E3779B9u; }   // LOP3 + IADD
      if (OP == 5) { a[k] = a[k] ^ b[k] ^ 0x12345u; }   // LOP3 only
    }
  }
  uint32_t t = 0;
#pragma unroll
  for (int k = 0; k < ILP; ++k) t += a[k] + b[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}
template <int OP>
void run(const char* name, uint32_t* d, int sms) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = sms * 8;
  probe<OP><<<grid, 256>>>(d, 1, 16);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  probe<OP><<<grid, 256>>>(d, 1, ITER);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double it = (double)grid * 256 * ITER * ILP;
  printf("%-28s %8.3f ms  = %6.1f iter/clk/SM @1.965GHz\n", name, ms, it / (ms * 1e-3) / sms / 1.965e9);
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  uint32_t* d; cudaMalloc(&d, 4 * p.multiProcessorCount * 8 * 256);
  run<0>("mul.wide.u32 + xor", d, p.multiProcessorCount);
  run<1>("mul.hi + mul.lo + xor", d, p.multiProcessorCount);
  run<2>("mul.lo + xor", d, p.multiProcessorCount);
  run<3>("mul.hi + xor", d, p.multiProcessorCount);
  run<4>("xor + add", d, p.multiProcessorCount);
  run<5>("xor3", d, p.multiProcessorCount);
  return 0;
}
