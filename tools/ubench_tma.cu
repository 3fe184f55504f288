// ubench_tma.cu -- per-SM throughput of 1-D bulk copies (cp.async.bulk, SASS UBLKCP) global/L2 -> shared memory:
// every CTA (one per SM) streams `chunk`-byte pieces of an L2-resident (or HBM-sized) buffer into a ring of `depth`
// shared-memory slots; a consumer thread frees a slot as soon as it is full.  Development tool (sizes the pipelines
// of k_generate.cuh / k_dense.cuh), not product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/ubench_tma tools/ubench_tma.cu && tools/ubench_tma
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(
          smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// split: number of separate bulk copies a chunk is issued as
__global__ void __launch_bounds__(64, 1) stream_kernel(const uint8_t* __restrict__ src, size_t span, uint32_t chunk, uint32_t depth,
                                                       uint32_t split, uint32_t iters, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t ring[];
  __shared__ uint64_t full[16], empty[16];
  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < depth; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(full + s)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(empty + s)));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long long t0 = clock64();
  if (threadIdx.x == 0) {          // producer
    size_t off = ((size_t)blockIdx.x * 7919u * chunk) % span;
    for (uint32_t i = 0; i < iters; ++i) {
      const uint32_t s = i % depth, r = i / depth;
      if (r) mbar_wait(empty + s, (r - 1) & 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(full + s)), "r"(chunk) : "memory");
      const uint32_t piece = chunk / split;
      for (uint32_t k = 0; k < split; ++k)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(ring + (size_t)s * chunk + k * piece)),
                     "l"(src + off + k * piece), "r"(piece), "r"(smem_u32(full + s))
                     : "memory");
      off += chunk;
      if (off + chunk > span) off = 0;
    }
  } else if (threadIdx.x == 32) {  // consumer: frees the slot at once
    for (uint32_t i = 0; i < iters; ++i) {
      const uint32_t s = i % depth, r = i / depth;
      mbar_wait(full + s, r & 1);
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(empty + s)) : "memory");
    }
    cycles[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  int dev = 0, sms = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const size_t big = (size_t)4 << 30;
  uint8_t* buf;
  CK(cudaMalloc(&buf, big));
  CK(cudaMemset(buf, 1, big));
  long long* dcyc;
  CK(cudaMalloc(&dcyc, sms * 8));
  printf("%d SMs; per-SM bytes/clk of cp.async.bulk global->shared (all SMs streaming)\n", sms);
  printf("%-10s %-8s %-6s %-6s %10s %12s\n", "source", "chunk", "depth", "split", "B/clk/SM", "chip TB/s@1.965");
  for (int src = 0; src < 2; ++src) {
    const size_t span = src == 0 ? ((size_t)48 << 20) : big;   // 48 MB stays in L2; 4 GiB streams from HBM
    for (uint32_t chunk : {8192u, 16384u, 32768u, 65536u}) {
      for (uint32_t total : {65536u, 131072u, 196608u}) {
        const uint32_t depth = total / chunk;
        if (depth < 1 || depth > 16) continue;
        for (uint32_t split : {1u, 4u}) {
          if (chunk / split < 2048) continue;
          const uint32_t iters = (uint32_t)(((size_t)64 << 20) / chunk);   // 64 MB per SM
          CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)total));
          stream_kernel<<<sms, 64, total>>>(buf, span, chunk, depth, split, iters, dcyc);   // warm (L2 fill)
          stream_kernel<<<sms, 64, total>>>(buf, span, chunk, depth, split, iters, dcyc);
          CK(cudaDeviceSynchronize());
          long long h[256];
          CK(cudaMemcpy(h, dcyc, sms * 8, cudaMemcpyDeviceToHost));
          double mx = 0;
          for (int i = 0; i < sms; ++i) mx = h[i] > mx ? (double)h[i] : mx;
          const double bpc = (double)iters * chunk / mx;
          printf("%-10s %-8u %-6u %-6u %10.1f %12.2f\n", src == 0 ? "L2 48MB" : "HBM 4GiB", chunk, depth, split, bpc, bpc * sms * 1.965e9 / 1e12);
        }
      }
    }
  }
  return 0;
}
