// ubench_rf.cu -- does operand bandwidth (register-file reads) cap the issue rate of an FP64 + integer mix on sm_100a?
// Each trip: NF FP64 FMAs whose three 64-bit sources are all registers (6 x 32-bit reads each) or register + 2 immediates,
// co-issued with NI integer ops that read 3 registers (IMAD / LOP3) -- the shape of generate's inner loop.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -fmad=false -o ubench_rf ubench_rf.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int ITER = 2048;

template <int MODE>
__global__ void __launch_bounds__(256) probe(double* out, uint32_t seed, int iters) {
  double a[8], b[8], c[8]; uint32_t n[8], m[8], q[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    a[k] = 1.0 + 1e-9 * (threadIdx.x + k + seed); b[k] = 1.0 + 1e-12 * (k + seed); c[k] = 1e-13 * (threadIdx.x + 3 * k);
    n[k] = threadIdx.x * 7u + k + seed; m[k] = n[k] * 3u + 1u; q[k] = seed + k;
  }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      // FP64 part
      if (MODE == 0 || MODE == 2 || MODE == 4) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[k]) : "d"(1.000001), "d"(0.5));          // 1 register source
      if (MODE == 1 || MODE == 3 || MODE == 5) asm volatile("fma.rn.f64 %0, %1, %0, %2;" : "+d"(a[k]) : "d"(b[k]), "d"(c[(k + 1) & 7]));   // 3 register sources
      // integer part
      if (MODE == 2 || MODE == 3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(n[k]) : "r"(m[k]), "r"(q[(k + 3) & 7]));            // LOP3, 3 register sources
      if (MODE == 4 || MODE == 5) {
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(n[k]) : "r"(m[k]), "r"(q[(k + 3) & 7]));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(m[k]) : "r"(q[k]));                                                                      // + IADD
      }
    }
  }
  double s = 0; uint32_t t = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) { s += a[k] + b[k] + c[k]; t += n[k] + m[k] + q[k]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + t;
}

template <int MODE>
void run(const char* name, double* d, int sms, int warps_per_sched) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int ctas_per_sm = warps_per_sched / 2;          // 256 threads = 8 warps = 2 per scheduler
  const int grid = sms * ctas_per_sm;
  probe<MODE><<<grid, 256>>>(d, 1, 16);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  probe<MODE><<<grid, 256>>>(d, 1, ITER);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double cyc = ms * 1e-3 * 1.965e9 / (warps_per_sched * (double)ITER);
  printf("%-58s %2d warps/sched  %6.2f scheduler cycles per trip\n", name, warps_per_sched, cyc);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  double* d; cudaMalloc(&d, sizeof(double) * p.multiProcessorCount * 8 * 256);
  const int S = p.multiProcessorCount;
  printf("%s, %d SMs; 8 FMAs alone = 17 cycles per trip\n", p.name, S);
  for (int w : {4, 8}) {
    run<0>("8 DFMA (1 register source)", d, S, w);
    run<1>("8 DFMA (3 register sources)", d, S, w);
    run<2>("8 DFMA (1 reg) + 8 LOP3 (3 regs)", d, S, w);
    run<3>("8 DFMA (3 regs) + 8 LOP3 (3 regs)", d, S, w);
    run<4>("8 DFMA (1 reg) + 8 LOP3 + 8 IADD", d, S, w);
    run<5>("8 DFMA (3 regs) + 8 LOP3 + 8 IADD", d, S, w);
  }
  return 0;
}
