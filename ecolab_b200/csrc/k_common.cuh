// k_common.cuh -- device-side layout, RNG and rounding shared by every kernel.
//
// Density layout in HBM ("blocked species-major", tile width TW in {4,8,16,32} cells):
// a grid row y of the slab is split into tiles of TW consecutive cells; a tile is the
// contiguous block  int32 [cap species rows][TW cells]; tiles of one grid row are
// contiguous, grid rows follow each other, one ghost row above and below the slab:
//     addr(y, x, s) = ((y*ntx + x/TW)*cap + s)*TW + x%TW          (ints, y relative to the slab)
// Consequences: a tile is ONE contiguous bulk (TMA) copy into shared memory for
// `generate`; a warp access = 128/TW species rows x TW cells = 512 contiguous bytes as
// 16-byte vectors (each lane owns 4 consecutive cells of one species row); a grid row is
// one contiguous halo message; adding/removing a species is a row operation inside
// every tile.  TW is chosen from the species capacity so that three tiles fit in one
// SM's shared memory (three resident generate CTAs overlap each other's loads).
//
// RNG contract (shared with oracle/ecolab_oracle.cc): counter-based Philox4x32-10,
//     block(x, y, slot, t, stage) = philox(ctr = {(x>>2) | stage<<30, y, slot, t_lo}, key = {seed, t_hi})
//     draw(cell (x,y), slot, t, stage) = block[x & 3]
// i.e. one Philox call yields the draws of 4 consecutive cells of one species row --
// exactly the 4 cells a lane owns.  stage: 0 generate, 1 mutate, 2 salt, 3 synthetic fill.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace eco {

constexpr unsigned FULL = 0xffffffffu;
constexpr double TWO52 = 4503599627370496.0;

struct Layout {
  int32_t* base;          // owned row 0 of the buffer (ghost row -1 lies row_stride before it)
  size_t tile_stride;     // cap * TW ints
  size_t row_stride;      // ntx * tile_stride ints
  uint32_t tw;            // tile width (cells)
  uint32_t ntx, nx, ny;   // tiles per grid row, grid width, OWNED rows
  uint32_t y0, nyg;       // first owned global row, global rows
  uint32_t nsp, cap;      // live species rows, allocated rows (multiple of 32, rows >= nsp are zero)
};

// exact for 0 <= n < 2^32, on the FP64 pipe (no I2F conversion-pipe instruction)
__device__ __forceinline__ double u2d(uint32_t n) {
  return __dadd_rn(__hiloint2double(0x43300000, (int)n), -TWO52);
}
// exact for any int32
__device__ __forceinline__ double i2d(int n) {
  return __dadd_rn(__hiloint2double(0x43300000, n ^ (int)0x80000000), -(TWO52 + 2147483648.0));
}

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                               uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned long long p0 = (unsigned long long)0xD2511F53u * c0;
    const unsigned long long p1 = (unsigned long long)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
// the four draws of cells 4*xq .. 4*xq+3 of grid row y, species slot `slot`, timestep t
__device__ __forceinline__ uint4 draw4(uint32_t xq, uint32_t y, uint32_t slot, unsigned long long t, uint32_t seed,
                                       uint32_t stage) {
  return philox4x32_10(xq | (stage << 30), y, slot, (uint32_t)t, seed, (uint32_t)(t >> 32));
}

// EcolabPoint::ROUND (reference models/ecolab_model.cc:57-66) with the draw u injected:
//   x<=0 -> 0 ; x >= INT_MAX-1 -> INT_MAX-1 ; else trunc(x) + (frac(x)*4294967295 > u)
// trunc and the comparison use directed-rounding adds against 2^52 (exact for 0 < x < 2^31), which keeps the
// whole function on the FP64 and integer pipes:  p > u  <=>  ceil(p) > u  for integer u.
__device__ __forceinline__ int round_stoch(double x, uint32_t u) {
  const double t = __dadd_rz(x, TWO52);                   // 2^52 + trunc(x)
  const int ti = __double2loint(t);
  const double f = __dadd_rn(x, -__dadd_rn(t, -TWO52));   // frac(x), exact
  const double p = __dmul_rn(f, 4294967295.0);
  const uint32_t cp = (uint32_t)__double2loint(__dadd_ru(p, TWO52));   // ceil(p)
  int r = ti + (cp > u ? 1 : 0);
  r = (x >= 2147483646.0) ? 2147483646 : r;
  r = (x <= 0.0) ? 0 : r;
  return r;
}

// the same for 0 <= x < 2^31-256 (callers check the range on x's high word): both clamps are no-ops there.
// x = +0 gives trunc 0, frac 0, ceil(0) = 0 > u false -> 0, as the reference's x<=0 branch.
__device__ __forceinline__ int round_stoch_inrange(double x, uint32_t u) {
  const double t = __dadd_rz(x, TWO52);
  const double f = __dadd_rn(x, -__dadd_rn(t, -TWO52));
  const uint32_t cp = (uint32_t)__double2loint(__dadd_ru(__dmul_rn(f, 4294967295.0), TWO52));
  return __double2loint(t) + (cp > u ? 1 : 0);
}

// ---- bounds-assert build (-DECO_BOUNDS; libecolab_b200_bounds.so, tests/test_gpu_bounds.py) ---------------------------
// compute-sanitizer cannot run on the GPU pool, so the same question -- does any kernel touch a byte outside the density
// buffers? -- is answered by a second build of the library in which every global access of the kernels that move densities
// is checked against the table of registered ranges (this rank's two buffers, the old buffer during a re-layout, the
// neighbour ranks' boundary rows).  A violation records its site and address; nothing is checked in the product build.
#ifdef ECO_BOUNDS
struct BoundsTable {
  const char* lo[6];
  const char* hi[6];
  unsigned int bad;            // number of violations
  unsigned int site;           // site code of the last one
  unsigned long long addr;     // its address
};
__device__ BoundsTable g_eco_bounds;
__device__ __forceinline__ void eco_chk(const void* ptr, size_t bytes, unsigned int site) {
  const char* c = static_cast<const char*>(ptr);
  bool ok = false;
#pragma unroll
  for (int i = 0; i < 6; ++i) ok |= g_eco_bounds.lo[i] != nullptr && c >= g_eco_bounds.lo[i] && c + bytes <= g_eco_bounds.hi[i];
  if (!ok) {
    atomicAdd(&g_eco_bounds.bad, 1u);
    g_eco_bounds.site = site;
    g_eco_bounds.addr = (unsigned long long)ptr;
  }
}
#define ECO_CHK(ptr, bytes, site) ::eco::eco_chk((ptr), (bytes), (site))
// negative control for the test: one access that is in range and one that straddles the end of the buffer
__global__ void bounds_selftest_kernel(const int32_t* buf, size_t ints) {
  ECO_CHK(buf + ints - 4, 16, 90u);   // the last 16 bytes: fine
  ECO_CHK(buf + ints - 3, 16, 91u);   // 4 bytes past the end: must be reported
}
#else
#define ECO_CHK(ptr, bytes, site) ((void)0)
#endif

// ---- mbarrier / bulk-copy (TMA unit) wrappers ------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

// lane -> (species row inside a warp-op, quad of cells) for tile width TW
template <int TW>
struct LaneMap {
  static constexpr int QPR = TW / 4;        // quads (lanes) per species row
  static constexpr int ROWS = 32 / QPR;     // species rows per warp-op
  __device__ static __forceinline__ uint32_t row(uint32_t lane) { return lane / QPR; }
  __device__ static __forceinline__ uint32_t quad(uint32_t lane) { return lane % QPR; }
};

}  // namespace eco
