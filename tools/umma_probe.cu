// umma_probe.cu -- one-CTA probe of tcgen05.mma kind::i8 (and kind::f16) with un-swizzled K-major shared-memory
// operands: pins the meaning of the shared-memory descriptor's LBO / SBO fields and of the instruction descriptor
// before the dense generate kernel (ecolab_b200/csrc/k_dense.cuh) relies on them.  Development tool, not product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_probe tools/umma_probe.cu && tools/umma_probe
#include <cuda_runtime.h>
#include <cuda_bf16.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout_type = 0) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) |
         (1ull << 46) | ((uint64_t)layout_type << 61);
}

// KIND 0: i8 (A signed, B unsigned, D s32)   KIND 1: bf16 x bf16 -> f32
template <int KIND>
__global__ void __launch_bounds__(128, 1) probe_kernel(const uint8_t* __restrict__ a_img, const uint8_t* __restrict__ b_img,
                                                       uint32_t* __restrict__ out, uint32_t kbytes, uint32_t lbo_a, uint32_t sbo_a,
                                                       uint32_t lbo_b, uint32_t sbo_b, uint32_t idesc, uint32_t N, uint32_t sw128,
                                                       uint32_t nrep, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t s_tmem;
  uint8_t* sa = smem;
  uint8_t* sb = smem + 128 * kbytes;
  for (uint32_t i = threadIdx.x; i < 128 * kbytes / 16; i += 128) reinterpret_cast<uint4*>(sa)[i] = reinterpret_cast<const uint4*>(a_img)[i];
  for (uint32_t i = threadIdx.x; i < N * kbytes / 16; i += 128) reinterpret_cast<uint4*>(sb)[i] = reinterpret_cast<const uint4*>(b_img)[i];
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the tensor core
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(256u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = s_tmem;
  if (threadIdx.x == 0) {
    // descriptors of the 4 K steps, built once: the timed loop is nothing but the MMA issue
    uint64_t ad[4], bd[4];
#pragma unroll
    for (uint32_t k = 0; k < 4; ++k) {
      // un-swizzled: 2 k16 chunks per MMA, a chunk is rows*16 bytes away; 128B swizzle: K runs inside the 128-byte row
      ad[k] = sw128 ? make_desc(smem_u32(sa) + k * 32, 16, 1024, 2) : make_desc(smem_u32(sa) + k * 2 * 128 * 16, lbo_a, sbo_a);
      bd[k] = sw128 ? make_desc(smem_u32(sb) + k * 32, 16, 1024, 2) : make_desc(smem_u32(sb) + k * 2 * N * 16, lbo_b, sbo_b);
    }
    const long long t0 = clock64();
    for (uint32_t rep = 0; rep < nrep; ++rep) {
#pragma unroll
      for (uint32_t k = 0; k < 4; ++k) {
        const uint32_t acc = (k || rep) ? 1u : 0u;
        if (KIND == 0)
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
                       "l"(ad[k]), "l"(bd[k]), "r"(idesc), "r"(acc)
                       : "memory");
        else
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
                       "l"(ad[k]), "l"(bd[k]), "r"(idesc), "r"(acc)
                       : "memory");
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t ok = 0;
    for (uint32_t spin = 0; spin < (1u << 24) && !ok; ++spin)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    if (cycles) *cycles = clock64() - t0;
  }
  // bounded wait: a wrong descriptor must not hang the box
  uint32_t done = 0;
  for (uint32_t spin = 0; spin < (1u << 22) && !done; ++spin) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done)
                 : "r"(smem_u32(&bar)), "r"(0u)
                 : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t row = warp * 32 + lane;
  if (!done) {
    out[(size_t)row * N] = 0xDEADBEEFu;
  } else {
    for (uint32_t c = 0; c < N; c += 4) {
      uint32_t r0, r1, r2, r3;
      const uint32_t taddr = tmem + ((warp * 32u) << 16) + c;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      out[(size_t)row * N + c] = r0; out[(size_t)row * N + c + 1] = r1;
      out[(size_t)row * N + c + 2] = r2; out[(size_t)row * N + c + 3] = r3;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u));
}

// canonical un-swizzled K-major image: [k16][row group of 8][8 rows][16 bytes]
static bool g_sw128 = false;
static size_t img_off(uint32_t rows, uint32_t r, uint32_t kbyte) {
  // 128-byte swizzle (Swizzle<3,4,3>): atoms of 8 rows x 128 B, 16-byte chunk index XOR row-in-atom; K extent = 128 B
  if (g_sw128) return (size_t)(r / 8) * 1024 + (r % 8) * 128 + (((kbyte / 16) ^ (r % 8)) * 16) + kbyte % 16;
  return (size_t)(kbyte / 16) * rows * 16 + (size_t)(r / 8) * 128 + (r % 8) * 16 + kbyte % 16;
}

int main() {
  const uint32_t M = 128, KB = 128;
  int fails = 0;
  for (int sw = 0; sw < 2; ++sw)
  for (uint32_t N : {128u, 64u, 256u}) {
    g_sw128 = sw != 0;
    for (int kind = 0; kind < 2; ++kind) {
      std::vector<uint8_t> a(M * KB), b(N * KB);
      std::vector<int8_t> A(M * KB);
      std::vector<uint8_t> B(N * KB);
      std::vector<float> Af(M * KB / 2), Bf(N * KB / 2);
      srand(1234 + N + kind);
      if (kind == 0) {
        for (uint32_t r = 0; r < M; ++r) for (uint32_t k = 0; k < KB; ++k) { A[r * KB + k] = (int8_t)(rand() % 256 - 128); a[img_off(M, r, k)] = (uint8_t)A[r * KB + k]; }
        for (uint32_t r = 0; r < N; ++r) for (uint32_t k = 0; k < KB; ++k) { B[r * KB + k] = (uint8_t)(rand() % 256); b[img_off(N, r, k)] = B[r * KB + k]; }
      } else {
        for (uint32_t r = 0; r < M; ++r) for (uint32_t k = 0; k < KB / 2; ++k) {
          const float v = (float)(rand() % 17 - 8); Af[r * KB / 2 + k] = v;
          __nv_bfloat16 h = __float2bfloat16(v); memcpy(&a[img_off(M, r, 2 * k)], &h, 2);
        }
        for (uint32_t r = 0; r < N; ++r) for (uint32_t k = 0; k < KB / 2; ++k) {
          const float v = (float)(rand() % 9 - 4); Bf[r * KB / 2 + k] = v;
          __nv_bfloat16 h = __float2bfloat16(v); memcpy(&b[img_off(N, r, 2 * k)], &h, 2);
        }
      }
      uint8_t *da, *db; uint32_t* dout;
      CK(cudaMalloc(&da, a.size())); CK(cudaMalloc(&db, b.size())); CK(cudaMalloc(&dout, M * N * 4));
      CK(cudaMemcpy(da, a.data(), a.size(), cudaMemcpyHostToDevice));
      CK(cudaMemcpy(db, b.data(), b.size(), cudaMemcpyHostToDevice));
      // idesc: c_format [4,6) (1 = f32, 2 = s32), a_format [7,10), b_format [10,13), n>>3 [17,23), m>>4 [24,29)
      const uint32_t idesc = kind == 0 ? ((2u << 4) | (1u << 7) | (0u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24))
                                       : ((1u << 4) | (1u << 7) | (1u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24));
      // (the swapped reading, LBO = row-group stride / SBO = k16 stride, raises an illegal memory access on B200: measured once, removed)
      struct V { const char* name; uint32_t la, sa, lb, sb; } variants[] = {{"LBO=k16 stride, SBO=row-group stride", M * 16, 128, N * 16, 128}};
      for (auto& v : variants) {
        CK(cudaMemset(dout, 0xFF, M * N * 4));
        const size_t smem = (size_t)(M + N) * KB;
        auto kern = kind == 0 ? probe_kernel<0> : probe_kernel<1>;
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        long long* dcyc; CK(cudaMalloc(&dcyc, 8)); CK(cudaMemset(dcyc, 0, 8));
        kern<<<1, 128, smem>>>(da, db, dout, KB, v.la, v.sa, v.lb, v.sb, idesc, N, (uint32_t)sw, 1u, dcyc);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("kind %d N %u variant '%s': CUDA error %s\n", kind, N, v.name, cudaGetErrorString(e)); return 3; }
        std::vector<uint32_t> out(M * N);
        CK(cudaMemcpy(out.data(), dout, M * N * 4, cudaMemcpyDeviceToHost));
        size_t bad = 0;
        bool timeout = out[0] == 0xDEADBEEFu;
        for (uint32_t i = 0; i < M && !timeout; ++i)
          for (uint32_t j = 0; j < N; ++j) {
            if (kind == 0) {
              int32_t ref = 0;
              for (uint32_t k = 0; k < KB; ++k) ref += (int32_t)A[i * KB + k] * (int32_t)B[j * KB + k];
              bad += (int32_t)out[i * N + j] != ref;
            } else {
              float ref = 0;
              for (uint32_t k = 0; k < KB / 2; ++k) ref += Af[i * KB / 2 + k] * Bf[j * KB / 2 + k];
              float got; memcpy(&got, &out[i * N + j], 4);
              bad += got != ref;
            }
          }
        // throughput: 512 back-to-back repetitions of the 4 k-steps on the same operands
        kern<<<1, 128, smem>>>(da, db, dout, KB, v.la, v.sa, v.lb, v.sb, idesc, N, (uint32_t)sw, 512u, dcyc);
        CK(cudaDeviceSynchronize());
        long long cyc = 0; CK(cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost)); cudaFree(dcyc);
        printf("kind %s N %3u layout %s: %s (%zu / %u mismatches)  %.1f cycles per MMA (M128 x N%u x 32 B)\n", kind == 0 ? "i8 " : "f16", N,
               sw ? "SWIZZLE_128B" : "un-swizzled ", timeout ? "TIMEOUT" : bad ? "WRONG" : "OK", bad, M * N, (double)cyc / (512.0 * 4), N);
        if (&v == &variants[0] && (timeout || bad)) ++fails;
      }
      cudaFree(da); cudaFree(db); cudaFree(dout);
    }
  }
  printf(fails ? "PROBE: expected variant FAILED in %d case(s)\n" : "PROBE: expected variant OK everywhere\n", fails);
  return 0;
}
