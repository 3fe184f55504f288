// ubench_i8peak.cu -- the tensor-core ceiling the dense generate kernel (ecolab_b200/csrc/k_dense.cuh) is measured against:
// every SM issues back-to-back tcgen05.mma.cta_group::1.kind::i8 (M128 x N x K32 B, SWIZZLE_128B operands resident in
// shared memory, no loads, no epilogue) for ~0.15 s per pass; reported per case: cycles per MMA, the SM clock the
// kernel really ran at (cycles / globaltimer ns -- the int8 pipe is power-capped), and the whole-GPU rate in TOP/s
// (2 ops per MAC).  Operand bytes: random (the dense path's regime) or zero (no datapath toggling: shows what the power cap costs).
// Development tool, not product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/ubench_i8peak tools/ubench_i8peak.cu && tools/ubench_i8peak
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// SWIZZLE_128B K-major descriptor: SBO = 1024 (8 rows x 128 B), version bit 46, layout type 2
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

// NACC accumulators of N columns are cycled so that consecutive MMAs do not depend on each other through one accumulator
__global__ void __launch_bounds__(128, 1) peak_kernel(const uint8_t* __restrict__ a_img, const uint8_t* __restrict__ b_img, uint32_t N,
                                                      uint32_t nacc, uint32_t nrep, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t s_tmem;
  uint8_t* sa = smem;                 // 128 rows x 128 B
  uint8_t* sb = smem + 128 * 128;     // N rows x 128 B
  for (uint32_t i = threadIdx.x; i < 128 * 128 / 16; i += 128) reinterpret_cast<uint4*>(sa)[i] = reinterpret_cast<const uint4*>(a_img)[i];
  for (uint32_t i = threadIdx.x; i < N * 128 / 16; i += 128) reinterpret_cast<uint4*>(sb)[i] = reinterpret_cast<const uint4*>(b_img)[i];
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = s_tmem;
  if (threadIdx.x == 0) {
    uint64_t ad[4], bd[4];
#pragma unroll
    for (uint32_t k = 0; k < 4; ++k) { ad[k] = make_desc(smem_u32(sa) + k * 32); bd[k] = make_desc(smem_u32(sb) + k * 32); }
    // idesc: D s32, A s8, B u8, n>>3 at [17,23), m>>4 at [24,29)
    const uint32_t idesc = (2u << 4) | (1u << 7) | (0u << 10) | ((N >> 3) << 17) | ((128u >> 4) << 24);
    unsigned long long ns0, ns1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns0));
    const long long t0 = clock64();
    uint32_t a = 0;
    for (uint32_t rep = 0; rep < nrep; ++rep) {
#pragma unroll
      for (uint32_t k = 0; k < 4; ++k) {
        const uint32_t acc = rep >= nacc ? 1u : (k ? 1u : 0u);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem + a * N),
                     "l"(ad[k]), "l"(bd[k]), "r"(idesc), "r"(acc)
                     : "memory");
      }
      if (++a == nacc) a = 0;
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t ok = 0;
    for (uint32_t spin = 0; spin < (1u << 30) && !ok; ++spin)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    const long long t1 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
    out[blockIdx.x * 4 + 0] = (unsigned long long)(t1 - t0);
    out[blockIdx.x * 4 + 1] = ns1 - ns0;
    out[blockIdx.x * 4 + 2] = ok;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u));
}

int main() {
  int nsm = 0;
  CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0));
  uint8_t *da, *db;
  unsigned long long* dout;
  CK(cudaMalloc(&da, 128 * 128)); CK(cudaMalloc(&db, 256 * 128)); CK(cudaMalloc(&dout, nsm * 4 * 8));
  const size_t smem = (size_t)(128 + 256) * 128;
  CK(cudaFuncSetAttribute(peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  printf("%d SMs; one CTA per SM, one issuing thread; 2 ops per MAC\n", nsm);
  for (int data = 0; data < 2; ++data) {
    std::vector<uint8_t> a(128 * 128), b(256 * 128);
    srand(7);
    for (auto& v : a) v = data ? (uint8_t)(rand() & 255) : 0;
    for (auto& v : b) v = data ? (uint8_t)(rand() & 255) : 0;
    CK(cudaMemcpy(da, a.data(), a.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(db, b.data(), b.size(), cudaMemcpyHostToDevice));
    for (uint32_t N : {128u, 256u}) {
      const uint32_t nacc = 512 / N >= 3 ? 3 : 512 / N;       // 3 accumulators at N = 128 (the dense path's shape), 2 at N = 256
      const uint32_t nrep = 480000u * (256u / N);             // ~ 2.5 x 10^8 cycles at the 64 / 128-cycle floor: ~140 ms per pass
      for (int pass = 0; pass < 4; ++pass) {                  // pass 0 = warm-up; 1-3 reported (the clock settles under the power cap)
        CK(cudaMemset(dout, 0, nsm * 4 * 8));
        CK(cudaEventRecord(e0));
        peak_kernel<<<nsm, 128, smem>>>(da, db, N, nacc, nrep, dout);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        if (!pass) continue;
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        std::vector<unsigned long long> o(nsm * 4);
        CK(cudaMemcpy(o.data(), dout, o.size() * 8, cudaMemcpyDeviceToHost));
        double cyc = 0, ns = 0; int bad = 0;
        for (int i = 0; i < nsm; ++i) { cyc += (double)o[i * 4]; ns += (double)o[i * 4 + 1]; bad += o[i * 4 + 2] != 1; }
        cyc /= nsm; ns /= nsm;
        const double macs = (double)nsm * nrep * 4.0 * 128.0 * N * 32.0;
        printf("operands %-6s N %3u: %6.1f cycles per MMA, %7.3f ms, SM clock %4.0f MHz in the kernel, %7.1f TOP/s int8%s\n",
               data ? "random" : "zero", N, cyc / (nrep * 4.0), ms, cyc / ns * 1e3, 2.0 * macs / (ms * 1e-3) * 1e-12, bad ? "  (TIMEOUT)" : "");
      }
    }
  }
  return 0;
}
