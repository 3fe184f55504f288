// ubench_pipes.cu -- instruction-throughput probes on sm_100a for the pipes the generate kernel leans on:
// FP64 add/mul, int->double conversion (I2F.F64), double->int (F2I), IMAD.WIDE (Philox), DSETP.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -fmad=false -o ubench_pipes ubench_pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITER = 4096;
constexpr int ILP = 8;

template <int OP>
__global__ void __launch_bounds__(256) probe(double* out, uint32_t seed, int iters) {
  double a[ILP]; uint32_t n[ILP]; unsigned long long w[ILP];
#pragma unroll
  for (int k = 0; k < ILP; ++k) { a[k] = 1.0 + 1e-9 * (threadIdx.x + k + seed); n[k] = threadIdx.x * 7u + k + seed; w[k] = n[k]; }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < ILP; ++k) {
      if (OP == 0) a[k] = __dadd_rn(a[k], 1.000001);
      if (OP == 1) a[k] = __dmul_rn(a[k], 1.000001);
      if (OP == 2) { a[k] = __dadd_rn(a[k], (double)n[k]); n[k] += 3u; }            // I2F.F64.U32 + DADD + IADD
      if (OP == 3) { n[k] += (uint32_t)__double2int_rz(a[k]); a[k] = __dadd_rn(a[k], 0.5); }   // F2I.S32.F64 + DADD + IADD
      if (OP == 4) { w[k] = (unsigned long long)0xD2511F53u * (uint32_t)w[k] + (w[k] >> 32); }   // IMAD.WIDE
      if (OP == 5) { n[k] += (a[k] > 1.5) ? 1u : 0u; a[k] = __dadd_rn(a[k], 1e-7); }  // DSETP + DADD
      if (OP == 6) { a[k] = __dadd_rn(__hiloint2double(0x43300000, (int)n[k]), a[k]); n[k] += 3u; }   // magic-number convert folded in add
      if (OP == 7) { a[k] = __dadd_rz(a[k], 1.000001); }
      if (OP == 8) { a[k] = __dadd_rn(a[k], (double)(int)n[k]); n[k] += 3u; }          // I2F.F64.S32
    }
  }
  double s = 0; uint32_t t = 0;
#pragma unroll
  for (int k = 0; k < ILP; ++k) { s += a[k]; t += n[k] + (uint32_t)w[k]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + t;
}

template <int OP>
void run(const char* name, double ops_per_iter, double* d, int sms) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = sms * 8;
  probe<OP><<<grid, 256>>>(d, 1, 16);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  probe<OP><<<grid, 256>>>(d, 1, ITER);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double thread_iters = (double)grid * 256 * ITER * ILP;
  printf("%-34s %8.3f ms  %7.2f G thread-iter/s  = %6.1f iter/clk/SM @1.965GHz (%g instr of interest per iter)\n", name, ms,
         thread_iters / ms / 1e6, thread_iters / (ms * 1e-3) / sms / 1.965e9, ops_per_iter);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  double* d; cudaMalloc(&d, sizeof(double) * p.multiProcessorCount * 8 * 256);
  printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
  run<0>("DADD", 1, d, p.multiProcessorCount);
  run<7>("DADD.RZ", 1, d, p.multiProcessorCount);
  run<1>("DMUL", 1, d, p.multiProcessorCount);
  run<2>("I2F.F64.U32 + DADD + IADD", 1, d, p.multiProcessorCount);
  run<8>("I2F.F64.S32 + DADD + IADD", 1, d, p.multiProcessorCount);
  run<6>("magic(hi=0x4330) + DADD + IADD", 1, d, p.multiProcessorCount);
  run<3>("F2I.S32.F64 + DADD + IADD", 1, d, p.multiProcessorCount);
  run<4>("IMAD.WIDE.U32 (+shift/add)", 1, d, p.multiProcessorCount);
  run<5>("DSETP + SEL/IADD + DADD", 1, d, p.multiProcessorCount);
  return 0;
}
