// ubench_coissue.cu -- does an FP64 instruction share its issue path with the integer pipes on sm_100a?
// Per loop trip and thread: NF independent FP64 ops + NI integer ops of one kind, 8 warps per scheduler.  If the two
// overlap, the trip costs max(FP64 time, integer time); if they share a port it costs the sum.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -fmad=false -o ubench_coissue ubench_coissue.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITER = 2048;

// KIND: 0 none, 1 IMAD (mad.lo), 2 LOP3 (xor), 3 IMAD.WIDE (mul.wide), 4 IADD3 (add), 5 mov via IMAD.MOV? (prmt) , 6 SHF
template <int NF, int NI, int KIND, int FOP>
__global__ void __launch_bounds__(256) probe(double* out, uint32_t seed, int iters) {
  double a[8]; uint32_t n[8]; unsigned long long w[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { a[k] = 1.0 + 1e-9 * (threadIdx.x + k + seed); n[k] = threadIdx.x * 7u + k + seed; w[k] = n[k]; }
  const uint32_t mulv = seed * 2654435761u + 12345u;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < NF; ++k) {
      if (FOP == 0) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(a[k & 7]) : "d"(1.000001));
      if (FOP == 1) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[k & 7]) : "d"(1.000001), "d"(0.5));
      if (FOP == 2) asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(a[k & 7]) : "d"(1.000001));
    }
#pragma unroll
    for (int k = 0; k < NI; ++k) {
      if (KIND == 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(n[k & 7]) : "r"(mulv), "r"(seed));
      if (KIND == 2) asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[k & 7]) : "r"(mulv));
      if (KIND == 3) asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w[k & 7]) : "r"((uint32_t)w[k & 7]), "r"(mulv));
      if (KIND == 4) asm volatile("add.u32 %0, %0, %1;" : "+r"(n[k & 7]) : "r"(mulv));
      if (KIND == 6) asm volatile("shf.l.wrap.b32 %0, %0, %0, %1;" : "+r"(n[k & 7]) : "r"(seed));
    }
  }
  double s = 0; uint32_t t = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) { s += a[k]; t += n[k] + (uint32_t)w[k] + (uint32_t)(w[k] >> 32); }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + t;
}

template <int NF, int NI, int KIND, int FOP>
void run(const char* name, double* d, int sms) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = sms * 8;
  probe<NF, NI, KIND, FOP><<<grid, 256>>>(d, 1, 16);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  probe<NF, NI, KIND, FOP><<<grid, 256>>>(d, 1, ITER);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  // scheduler cycles per loop trip of ONE warp: elapsed cycles * 4 schedulers / (warps per SM * trips)
  const double warps_per_sm = 8.0 * 256 / 32;
  const double cyc = ms * 1e-3 * 1.965e9 * 4 / (warps_per_sm * ITER);
  printf("%-44s %7.3f ms   %6.2f scheduler cycles per trip\n", name, ms, cyc);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  double* d; cudaMalloc(&d, sizeof(double) * p.multiProcessorCount * 8 * 256);
  const int S = p.multiProcessorCount;
  printf("%s, %d SMs (cycles at 1.965 GHz; 8 FP64 ops alone = 8 x 2 cycles if the FP64 pipe is half rate)\n", p.name, S);
  run<8, 0, 0, 0>("8 DADD", d, S);
  run<8, 0, 0, 1>("8 DFMA", d, S);
  run<8, 0, 0, 2>("8 DMUL", d, S);
  run<0, 8, 1, 0>("8 IMAD", d, S);
  run<0, 8, 2, 0>("8 LOP3", d, S);
  run<0, 8, 3, 0>("8 IMAD.WIDE", d, S);
  run<0, 8, 4, 0>("8 IADD", d, S);
  run<0, 8, 6, 0>("8 SHF", d, S);
  run<8, 8, 1, 0>("8 DADD + 8 IMAD", d, S);
  run<8, 16, 1, 0>("8 DADD + 16 IMAD", d, S);
  run<8, 8, 2, 0>("8 DADD + 8 LOP3", d, S);
  run<8, 16, 2, 0>("8 DADD + 16 LOP3", d, S);
  run<8, 8, 4, 0>("8 DADD + 8 IADD", d, S);
  run<8, 4, 3, 0>("8 DADD + 4 IMAD.WIDE", d, S);
  run<8, 8, 3, 0>("8 DADD + 8 IMAD.WIDE", d, S);
  run<8, 8, 1, 1>("8 DFMA + 8 IMAD", d, S);
  run<8, 8, 2, 1>("8 DFMA + 8 LOP3", d, S);
  run<8, 16, 6, 0>("8 DADD + 16 SHF", d, S);
  run<0, 16, 1, 0>("16 IMAD", d, S);
  run<0, 16, 2, 0>("16 LOP3", d, S);
  return 0;
}
