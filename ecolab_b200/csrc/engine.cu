// engine.cu -- device store and sm_100a kernels for the EcoLab population update.
//
// Density layout in HBM ("blocked species-major"): the grid row y of the slab is
// split into tiles of 32 consecutive cells (x = 32*tx + lane); a tile is the
// contiguous block  int32 [cap species rows][32 lanes]  and tiles of one grid
// row are contiguous, so
//     addr(y, x, s) = ((y*ntx + x/32)*cap + s)*32 + x%32 .
// Consequences: a tile is one contiguous bulk (TMA) copy into shared memory for
// `generate`; every warp access (fixed species, 32 cells) is one 128-byte line;
// a grid row is one contiguous halo message; adding/removing a species is a row
// operation inside each tile.  One ghost row above and below the slab receives
// the neighbour slabs' boundary rows (multi-GPU migrate).
//
// Reference semantics (cited per kernel) come from models/ecolab_model.cc of
// highperformancecoder/ecolab; this file shares no code with it.
#include "engine.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <stdexcept>

#define CK(x) ::eco::check_cuda((int)(x), #x, __FILE__, __LINE__)

namespace eco {

void check_cuda(int err, const char* what, const char* file, int line) {
  if (err != (int)cudaSuccess) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", err,
             cudaGetErrorString((cudaError_t)err), file, line, what);
    throw std::runtime_error(buf);
  }
}

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr double TWO52 = 4503599627370496.0;

// ---------------------------------------------------------------------------
// exact int <-> double helpers on the FP64 pipe (no I2F/F2I conversion pipe)
// ---------------------------------------------------------------------------
// exact for 0 <= n < 2^32
__device__ __forceinline__ double u2d(uint32_t n) {
  return __dadd_rn(__hiloint2double(0x43300000, (int)n), -TWO52);
}
// exact for any int32
__device__ __forceinline__ double i2d(int n) {
  return __dadd_rn(__hiloint2double(0x43300000, n ^ (int)0x80000000), -(TWO52 + 2147483648.0));
}

// Philox4x32-10; ctr = {cell, slot>>2, t_lo, t_hi}, key = {seed, stage} (DESIGN.md "RNG contract")
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

// EcolabPoint::ROUND, reference models/ecolab_model.cc:57-66, with the draw u injected.
//   x<=0 -> 0 ; x >= INT_MAX-1 -> INT_MAX-1 ; else trunc(x) + (frac(x)*4294967295 > u)
// trunc and the comparison are done with directed-rounding adds against 2^52,
// which is exact for 0 < x < 2^31 and keeps everything on the FP64/int pipes:
//   p > u  <=>  ceil(p) > u  for integer u.
__device__ __forceinline__ int round_stoch(double x, uint32_t u) {
  const double t = __dadd_rz(x, TWO52);                // 2^52 + trunc(x)   (valid for 0<=x<2^32)
  const int ti = __double2loint(t);
  const double f = __dadd_rn(x, -__dadd_rn(t, -TWO52));  // frac(x), exact
  const double p = __dmul_rn(f, 4294967295.0);
  const uint32_t cp = (uint32_t)__double2loint(__dadd_ru(p, TWO52));   // ceil(p)
  int r = ti + (cp > u ? 1 : 0);
  r = (x >= 2147483646.0) ? 2147483646 : r;
  r = (x <= 0.0) ? 0 : r;
  return r;
}

// ---------------------------------------------------------------------------
// mbarrier / bulk-copy (TMA) wrappers
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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

// ---------------------------------------------------------------------------
// generate  (reference EcolabPoint::generate, models/ecolab_model.cc:90-112)
// ---------------------------------------------------------------------------
struct GenParams {
  int32_t* dens;          // owned row 0 of the current buffer
  size_t tile_stride;     // ints
  uint32_t ntiles, ntx, nx, y0;
  uint32_t nsp, ngroups;
  const uint32_t* gptr;
  const double* gval;
  const uint32_t* gcol;
  const double* diag4;
  const double* rr4;
  uint32_t seed, niter;
  unsigned long long t0;
};

// One row group (4 consecutive species rows) for one cell (= lane).  All loads of
// beta are warp-uniform; all shared-memory gathers are conflict-free (32
// consecutive words).  FP order is exactly the reference's: ir = diag*n, then
// += val*n[col] in COO order, then n + n*(r + ir), each op individually rounded.
__device__ __forceinline__ void gen_group(const int32_t* __restrict__ tile, uint32_t lane, uint32_t g,
                                          const GenParams& p, uint32_t cell, unsigned long long t,
                                          int out[4]) {
  const uint32_t base = g * 128u + lane;
  const double d0 = u2d((uint32_t)tile[base]), d1 = u2d((uint32_t)tile[base + 32]);
  const double d2 = u2d((uint32_t)tile[base + 64]), d3 = u2d((uint32_t)tile[base + 96]);
  const double2* dg = reinterpret_cast<const double2*>(p.diag4 + 4u * g);
  const double2 dga = __ldg(dg), dgb = __ldg(dg + 1);
  double ir0 = __dmul_rn(dga.x, d0), ir1 = __dmul_rn(dga.y, d1);
  double ir2 = __dmul_rn(dgb.x, d2), ir3 = __dmul_rn(dgb.y, d3);
  uint32_t j = __ldg(p.gptr + g);
  const uint32_t je = __ldg(p.gptr + g + 1);
  for (; j < je; ++j) {
    const double2* vp = reinterpret_cast<const double2*>(p.gval + 4ull * j);
    const double2 va = __ldg(vp), vb = __ldg(vp + 1);
    const uint4 c = __ldg(reinterpret_cast<const uint4*>(p.gcol + 4ull * j));
    ir0 = __dadd_rn(ir0, __dmul_rn(va.x, u2d((uint32_t)tile[c.x + lane])));
    ir1 = __dadd_rn(ir1, __dmul_rn(va.y, u2d((uint32_t)tile[c.y + lane])));
    ir2 = __dadd_rn(ir2, __dmul_rn(vb.x, u2d((uint32_t)tile[c.z + lane])));
    ir3 = __dadd_rn(ir3, __dmul_rn(vb.y, u2d((uint32_t)tile[c.w + lane])));
  }
  const double2* rp = reinterpret_cast<const double2*>(p.rr4 + 4u * g);
  const double2 ra = __ldg(rp), rb = __ldg(rp + 1);
  const double x0 = __dadd_rn(d0, __dmul_rn(d0, __dadd_rn(ra.x, ir0)));
  const double x1 = __dadd_rn(d1, __dmul_rn(d1, __dadd_rn(ra.y, ir1)));
  const double x2 = __dadd_rn(d2, __dmul_rn(d2, __dadd_rn(rb.x, ir2)));
  const double x3 = __dadd_rn(d3, __dmul_rn(d3, __dadd_rn(rb.y, ir3)));
  const uint4 u = philox4x32_10(cell, g, (uint32_t)t, (uint32_t)(t >> 32), p.seed, 0u /*stage generate*/);
  out[0] = round_stoch(x0, u.x);
  out[1] = round_stoch(x1, u.y);
  out[2] = round_stoch(x2, u.z);
  out[3] = round_stoch(x3, u.w);
}

// Persistent CTAs; each iteration: bulk-copy one tile (all species of 32 cells)
// into shared memory, run `niter` Jacobi steps, write the tile back in place.
//  DB=false: niter==1, single smem buffer, results stored straight to HBM
//            (one 128-byte line per warp per species row).
//  DB=true : two smem buffers, any niter (HBM traffic amortised over niter).
template <int THREADS, bool DB>
__global__ void __launch_bounds__(THREADS, 1) generate_kernel(const GenParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const uint32_t nsp_pad = p.ngroups * 4u;
  const uint32_t tile_words = nsp_pad * 32u;
  const uint32_t tile_bytes = tile_words * 4u;
  int32_t* buf0 = reinterpret_cast<int32_t*>(smem_raw);
  int32_t* buf1 = buf0 + (DB ? tile_words : 0);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + (size_t)tile_bytes * (DB ? 2 : 1));

  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = THREADS / 32;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  uint32_t phase = 0;
  for (uint32_t tile_id = blockIdx.x; tile_id < p.ntiles; tile_id += gridDim.x) {
    int32_t* gt = p.dens + (size_t)tile_id * p.tile_stride;
    if (threadIdx.x == 0) {
      mbar_expect_tx(bar, tile_bytes);
      // chunked so a single request stays modest; all complete on the same barrier
      for (uint32_t off = 0; off < tile_bytes; off += 32768u) {
        const uint32_t n = min(32768u, tile_bytes - off);
        bulk_g2s(reinterpret_cast<unsigned char*>(buf0) + off, reinterpret_cast<const unsigned char*>(gt) + off, n, bar);
      }
      const uint32_t nxt = tile_id + gridDim.x;
      if (nxt < p.ntiles) bulk_prefetch_l2(p.dens + (size_t)nxt * p.tile_stride, tile_bytes);
    }
    mbar_wait(bar, phase);
    phase ^= 1u;

    const uint32_t yl = tile_id / p.ntx, tx = tile_id - yl * p.ntx;
    const uint32_t cell = (p.y0 + yl) * p.nx + tx * 32u + lane;

    if (!DB) {
      for (uint32_t g = warp; g < p.ngroups; g += nwarps) {
        int o[4];
        gen_group(buf0, lane, g, p, cell, p.t0, o);
        int32_t* dst = gt + g * 128u + lane;
#pragma unroll
        for (int r = 0; r < 4; ++r)
          if (4u * g + r < p.nsp) dst[32 * r] = o[r];
      }
      __syncthreads();   // everyone done reading buf0 before the next bulk copy lands
    } else {
      int32_t* src = buf0;
      int32_t* dst = buf1;
      for (uint32_t step = 0; step < p.niter; ++step) {
        for (uint32_t g = warp; g < p.ngroups; g += nwarps) {
          int o[4];
          gen_group(src, lane, g, p, cell, p.t0 + step, o);
          int32_t* d = dst + g * 128u + lane;
#pragma unroll
          for (int r = 0; r < 4; ++r) d[32 * r] = o[r];
        }
        __syncthreads();
        int32_t* tmp = src; src = dst; dst = tmp;
      }
      // write back rows < nsp (src holds the result)
      const uint32_t nvec = p.nsp * 8u;   // int4 per tile
      const int4* s4 = reinterpret_cast<const int4*>(src);
      int4* g4 = reinterpret_cast<int4*>(gt);
      for (uint32_t i = threadIdx.x; i < nvec; i += THREADS) g4[i] = s4[i];
      __syncthreads();
    }
  }
}

// x = n + n*(r + ir) of ONE generate sub-step, before ROUND (reference :99-102), for every owned
// cell; densities are not modified.  Test/diagnostic path for the "pre-rounding float update
// within 1e-6" criterion: one warp per (tile, row group), operands straight from global memory.
__global__ void __launch_bounds__(256) pre_rounding_kernel(const GenParams p, double* __restrict__ out) {
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  const unsigned long long nwork = (unsigned long long)p.ntiles * p.ngroups;
  for (unsigned long long w = (unsigned long long)blockIdx.x * wpb + (threadIdx.x >> 5); w < nwork;
       w += (unsigned long long)gridDim.x * wpb) {
    const uint32_t g = (uint32_t)(w % p.ngroups), tile_id = (uint32_t)(w / p.ngroups);
    const int32_t* tile = p.dens + (size_t)tile_id * p.tile_stride;
    const uint32_t yl = tile_id / p.ntx, tx = tile_id - yl * p.ntx, x = tx * 32u + lane;
    double d[4], ir[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      d[r] = (4u * g + r < p.nsp) ? i2d(tile[(4u * g + r) * 32u + lane]) : 0.0;
      ir[r] = __dmul_rn(p.diag4[4u * g + r], d[r]);
    }
    for (uint32_t j = p.gptr[g]; j < p.gptr[g + 1]; ++j) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
        ir[r] = __dadd_rn(ir[r], __dmul_rn(p.gval[4ull * j + r], i2d(tile[p.gcol[4ull * j + r] + lane])));
    }
    if (x < p.nx) {
      double* o = out + ((size_t)yl * p.nx + x) * p.nsp;
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (4u * g + r < p.nsp) o[4u * g + r] = __dadd_rn(d[r], __dmul_rn(d[r], __dadd_rn(p.rr4[4u * g + r], ir[r])));
    }
  }
}

// ---------------------------------------------------------------------------
// migrate  (reference SpatialModel::migrate, models/ecolab_model.cc:616-703)
//   delta = sum over existing W,E,N,S neighbours of (int)(cm[s] * (n_nbr - n_own))
//   n_new = n_own + delta  (synchronous: reads `src`, writes `dst`)
// ---------------------------------------------------------------------------
struct MigParams {
  const int32_t* src;
  int32_t* dst;
  const int32_t* top_row;      // row "y0-1" (ghost or peer), may be null if y0==0
  const int32_t* bottom_row;   // row "y0+ny"
  size_t tile_stride, row_stride;
  uint32_t ntx, nx, ny, y0, nyg, nsp;
  uint32_t nchunks, chunk;     // species chunks per tile
  const double* cm;
  unsigned long long* total;
};

__device__ __forceinline__ int mig_term(double c, int nb, int own) {
  return __double2int_rz(__dmul_rn(c, i2d(nb - own)));
}

__global__ void __launch_bounds__(256) migrate_kernel(const MigParams p) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t wpb = blockDim.x >> 5;
  const unsigned long long nwork = (unsigned long long)p.ntx * p.ny * p.nchunks;
  unsigned long long acc = 0;
  for (unsigned long long w = (unsigned long long)blockIdx.x * wpb + (threadIdx.x >> 5); w < nwork;
       w += (unsigned long long)gridDim.x * wpb) {
    const uint32_t ch = (uint32_t)(w % p.nchunks);
    const uint32_t tile = (uint32_t)(w / p.nchunks);
    const uint32_t yl = tile / p.ntx, tx = tile - yl * p.ntx;
    const uint32_t x = tx * 32u + lane, yg = p.y0 + yl;
    const bool valid = x < p.nx;
    const bool has_w = valid && x > 0, has_e = x + 1 < p.nx;
    const bool has_n = valid && yg > 0, has_s = valid && yg + 1 < p.nyg;
    const int32_t* own_t = p.src + (size_t)tile * p.tile_stride;
    const int32_t* n_t = (yl > 0) ? own_t - p.row_stride : p.top_row + (size_t)tx * p.tile_stride;
    const int32_t* s_t = (yl + 1 < p.ny) ? own_t + p.row_stride : p.bottom_row + (size_t)tx * p.tile_stride;
    // one predicated load fetches both tile-edge neighbours: lane 0 <- west tile's lane 31, lane 31 <- east tile's lane 0
    const bool edge_w = lane == 0 && tx > 0, edge_e = lane == 31 && has_e;
    const int32_t* e_t = edge_w ? own_t - p.tile_stride + 31 : own_t + p.tile_stride - 31;
    int32_t* dst_t = p.dst + (size_t)tile * p.tile_stride;
    const uint32_t s0 = ch * p.chunk, s1 = min(p.nsp, s0 + p.chunk);
#pragma unroll 4
    for (uint32_t s = s0; s < s1; ++s) {
      const uint32_t o = s * 32u + lane;
      const int own = own_t[o];
      const int nn = has_n ? n_t[o] : own;
      const int ss = has_s ? s_t[o] : own;
      int ee = (edge_w || edge_e) ? e_t[o] : own;
      int wv = __shfl_up_sync(FULL, own, 1), ev = __shfl_down_sync(FULL, own, 1);
      if (lane == 0) wv = ee;
      if (lane == 31) ev = ee;
      wv = has_w ? wv : own;
      ev = (valid && has_e) ? ev : own;
      const double c = __ldg(p.cm + s);
      const int d = mig_term(c, wv, own) + mig_term(c, ev, own) + mig_term(c, nn, own) + mig_term(c, ss, own);
      dst_t[o] = valid ? own + d : 0;
      acc += (unsigned long long)(d < 0 ? -d : d);
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
  if (lane == 0 && acc) atomicAdd(p.total, acc);
}

// ---------------------------------------------------------------------------
// condense helpers (reference SpatialModel::condense :171-196, EcolabPoint::condense :124-137)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) alive_kernel(const int32_t* dens, size_t tile_stride, uint32_t ntiles,
                                                    uint32_t nsp, uint8_t* mask) {
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int32_t* t = dens + (size_t)tile * tile_stride;
    for (uint32_t s = threadIdx.x >> 5; s < nsp; s += wpb) {
      const bool any = __any_sync(FULL, t[s * 32u + lane] != 0);
      if (any && lane == 0) mask[s] = 1;   // benign race: all writers store 1
    }
  }
}

// in-place, order-preserving row compaction inside each tile; map[s] = new index of kept row s
__global__ void __launch_bounds__(1024) compact_kernel(int32_t* dens, size_t tile_stride, uint32_t ntiles,
                                                       uint32_t nsp_old, uint32_t nkeep, const uint8_t* keep,
                                                       const uint32_t* map) {
  const uint32_t lane = threadIdx.x & 31u, r = threadIdx.x >> 5;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int32_t* t = dens + (size_t)tile * tile_stride;
    for (uint32_t c0 = 0; c0 < nsp_old; c0 += 32) {
      const uint32_t s = c0 + r;
      int v = 0;
      bool k = false;
      if (s < nsp_old) { v = t[s * 32u + lane]; k = keep[s] != 0; }
      __syncthreads();   // all rows of this chunk are in registers before any is overwritten
      if (k) t[map[s] * 32u + lane] = v;
    }
    __syncthreads();
    for (uint32_t s = nkeep + r; s < nsp_old; s += 32) t[s * 32u + lane] = 0;
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// mutate (cell side): reference EcolabPoint::mutate, models/ecolab_model.cc:278-314
//   k = ROUND(ms[s] * n[c][s]) with the stage-1 Philox stream; records (cell, slot, k>0)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) mutate_scan_kernel(const int32_t* dens, size_t tile_stride, uint32_t ntiles,
                                                          uint32_t ntx, uint32_t nx, uint32_t y0, uint32_t nsp,
                                                          const double* ms, uint32_t seed, unsigned long long t,
                                                          MutRecord* rec, unsigned long long rec_cap,
                                                          unsigned long long* count) {
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  const uint32_t ngroups = (nsp + 3) / 4;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int32_t* tp = dens + (size_t)tile * tile_stride;
    const uint32_t yl = tile / ntx, tx = tile - yl * ntx;
    const uint32_t x = tx * 32u + lane;
    const uint32_t cell = (y0 + yl) * nx + x;
    for (uint32_t g = threadIdx.x >> 5; g < ngroups; g += wpb) {
      int n[4];
      bool anypos = false;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const uint32_t s = 4 * g + r;
        n[r] = (s < nsp && x < nx) ? tp[s * 32u + lane] : 0;
        anypos |= n[r] > 0;
      }
      if (!__any_sync(FULL, anypos)) continue;
      const uint4 u = philox4x32_10(cell, g, (uint32_t)t, (uint32_t)(t >> 32), seed, 1u /*stage mutate*/);
      const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const uint32_t s = 4 * g + r;
        if (n[r] > 0) {
          const int k = round_stoch(__dmul_rn(__ldg(ms + s), u2d((uint32_t)n[r])), uu[r]);
          if (k > 0) {
            const unsigned long long i = atomicAdd(count, 1ull);
            if (i < rec_cap) rec[i] = MutRecord{cell, s, (uint32_t)k};
          }
        }
      }
    }
  }
}

__global__ void mutate_apply_kernel(int32_t* dens, size_t tile_stride, uint32_t ntx, uint32_t nx, uint32_t y0,
                                    const MutRecord* rec, uint32_t nrec) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nrec) return;
  const MutRecord r = rec[i];
  const uint32_t yg = r.cell / nx, x = r.cell - yg * nx;
  const size_t tile = (size_t)(yg - y0) * ntx + x / 32;
  dens[tile * tile_stride + (size_t)r.slot * 32u + (x & 31u)] -= (int)r.count;
}

// new species rows: zero everywhere, then 1 in the birth cell (reference :272-275)
__global__ void zero_rows_kernel(int32_t* dens, size_t tile_stride, uint32_t ntiles, uint32_t s0, uint32_t s1) {
  const uint32_t per_tile = (s1 - s0) * 32u;
  const unsigned long long n = (unsigned long long)ntiles * per_tile;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t tile = (uint32_t)(i / per_tile), o = (uint32_t)(i % per_tile);
    dens[(size_t)tile * tile_stride + (size_t)s0 * 32u + o] = 0;
  }
}
__global__ void birth_kernel(int32_t* dens, size_t tile_stride, uint32_t ntx, uint32_t nx, uint32_t y0, uint32_t ny,
                             uint32_t s0, const uint32_t* cells, uint32_t m) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint32_t c = cells[i];
  const uint32_t yg = c / nx, x = c - yg * nx;
  if (yg < y0 || yg >= y0 + ny) return;
  const size_t tile = (size_t)(yg - y0) * ntx + x / 32;
  dens[tile * tile_stride + (size_t)(s0 + i) * 32u + (x & 31u)] = 1;
}

// ---------------------------------------------------------------------------
// observables
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) nsp_cell_kernel(const int32_t* dens, size_t tile_stride, uint32_t ntiles,
                                                       uint32_t ntx, uint32_t nx, uint32_t nsp, uint32_t* out) {
  __shared__ uint32_t cnt[32];
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    if (threadIdx.x < 32) cnt[threadIdx.x] = 0;
    __syncthreads();
    const int32_t* t = dens + (size_t)tile * tile_stride;
    uint32_t c = 0;
    for (uint32_t s = threadIdx.x >> 5; s < nsp; s += wpb) c += t[s * 32u + lane] != 0;
    atomicAdd(&cnt[lane], c);
    __syncthreads();
    const uint32_t yl = tile / ntx, tx = tile - yl * ntx;
    if (threadIdx.x < 32 && tx * 32u + lane < nx) out[(size_t)yl * nx + tx * 32u + lane] = cnt[lane];
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) totals_kernel(const int32_t* dens, size_t tile_stride, uint32_t ntiles,
                                                     uint32_t nsp, long long* out) {
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  for (uint32_t s = blockIdx.x * wpb + (threadIdx.x >> 5); s < nsp; s += gridDim.x * wpb) {
    long long acc = 0;
    for (uint32_t tile = 0; tile < ntiles; ++tile) acc += dens[(size_t)tile * tile_stride + s * 32u + lane];
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
    if (lane == 0) out[s] = acc;
  }
}

// ---------------------------------------------------------------------------
// layout conversion: cell-major [cell][nsp] <-> blocked
// ---------------------------------------------------------------------------
template <bool TO_BLOCKED>
__global__ void __launch_bounds__(1024) transpose_kernel(int32_t* blocked, size_t tile_stride, uint32_t ntx,
                                                         uint32_t nx, uint32_t ny, uint32_t nsp, int32_t* cellmajor) {
  __shared__ int32_t sm[32][33];
  const uint32_t tx_ = threadIdx.x & 31u, ty_ = threadIdx.x >> 5;
  const uint32_t nsb = (nsp + 31) / 32;
  const unsigned long long nwork = (unsigned long long)ntx * ny * nsb;
  for (unsigned long long w = blockIdx.x; w < nwork; w += gridDim.x) {
    const uint32_t sb = (uint32_t)(w % nsb);
    const uint32_t tile = (uint32_t)(w / nsb);
    const uint32_t yl = tile / ntx, tx = tile - yl * ntx;
    int32_t* bt = blocked + (size_t)tile * tile_stride;
    if (TO_BLOCKED) {
      // read: cell = ty_, species = sb*32 + tx_ (coalesced along species)
      const uint32_t x = tx * 32u + ty_, s = sb * 32u + tx_;
      sm[ty_][tx_] = (x < nx && s < nsp) ? cellmajor[((size_t)yl * nx + x) * nsp + s] : 0;
      __syncthreads();
      const uint32_t s2 = sb * 32u + ty_;
      if (s2 < nsp) bt[s2 * 32u + tx_] = sm[tx_][ty_];
      __syncthreads();
    } else {
      const uint32_t s2 = sb * 32u + ty_;
      sm[ty_][tx_] = (s2 < nsp) ? bt[s2 * 32u + tx_] : 0;
      __syncthreads();
      const uint32_t x = tx * 32u + ty_, s = sb * 32u + tx_;
      if (x < nx && s < nsp) cellmajor[((size_t)yl * nx + x) * nsp + s] = sm[tx_][ty_];
      __syncthreads();
    }
  }
}

// copy rows [0,nrows) of every tile between buffers with different capacities
__global__ void relayout_kernel(const int32_t* src, size_t src_stride, int32_t* dst, size_t dst_stride,
                                uint32_t ntiles, uint32_t nrows, uint32_t dst_rows) {
  const uint32_t per_tile = dst_rows * 32u;
  const unsigned long long n = (unsigned long long)ntiles * per_tile;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t tile = (uint32_t)(i / per_tile), o = (uint32_t)(i % per_tile);
    dst[(size_t)tile * dst_stride + o] = (o < nrows * 32u) ? src[(size_t)tile * src_stride + o] : 0;
  }
}

__global__ void fill_kernel(int32_t* dens, size_t tile_stride, uint32_t ntiles, uint32_t ntx, uint32_t nx,
                            uint32_t nsp, int32_t v) {
  const uint32_t per_tile = nsp * 32u;
  const unsigned long long n = (unsigned long long)ntiles * per_tile;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t tile = (uint32_t)(i / per_tile), o = (uint32_t)(i % per_tile);
    const uint32_t tx = tile % ntx;
    dens[(size_t)tile * tile_stride + o] = (tx * 32u + (o & 31u) < nx) ? v : 0;
  }
}

// synthetic initial state: n = lo + draw % span with the stage-3 Philox stream (same contract as the hot path:
// ctr = {cell, slot>>2, t=0}, key = {seed, 3}); used by benchmarks and full-size property tests
__global__ void __launch_bounds__(256) fill_random_kernel(int32_t* dens, size_t tile_stride, uint32_t ntiles, uint32_t ntx,
                                                          uint32_t nx, uint32_t y0, uint32_t nsp, uint32_t seed, int32_t lo,
                                                          uint32_t span) {
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5, ngroups = (nsp + 3) / 4;
  const unsigned long long nwork = (unsigned long long)ntiles * ngroups;
  for (unsigned long long w = (unsigned long long)blockIdx.x * wpb + (threadIdx.x >> 5); w < nwork;
       w += (unsigned long long)gridDim.x * wpb) {
    const uint32_t g = (uint32_t)(w % ngroups), tile = (uint32_t)(w / ngroups);
    const uint32_t yl = tile / ntx, tx = tile - yl * ntx, x = tx * 32u + lane;
    const uint32_t cell = (y0 + yl) * nx + x;
    const uint4 u = philox4x32_10(cell, g, 0u, 0u, seed, 3u);
    const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
    int32_t* t = dens + (size_t)tile * tile_stride;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const uint32_t s = 4u * g + r;
      if (s < nsp) t[s * 32u + lane] = x < nx ? lo + (int32_t)(uu[r] % span) : 0;
    }
  }
}

inline uint32_t round_up4(uint32_t n) { return (n + 3u) & ~3u; }

}  // namespace

// ===========================================================================
// Engine
// ===========================================================================
#define ST ((cudaStream_t)stream_)

Engine::Engine(int device) : device_(device) {
  CK(cudaSetDevice(device_));
  cudaStream_t s;
  CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  stream_ = s;
  own_stream_ = true;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device_));
  num_sms_ = prop.multiProcessorCount;
  max_smem_ = (int)prop.sharedMemPerBlockOptin;
  CK(cudaMalloc(&d_total_, 4 * sizeof(unsigned long long)));
  CK(cudaMemset(d_total_, 0, 4 * sizeof(unsigned long long)));
  CK(cudaMallocHost(&h_total_, 4 * sizeof(unsigned long long)));
  cudaEvent_t ev;
  CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  ev_total_ = ev;
}

Engine::~Engine() {
  cudaSetDevice(device_);
  cudaDeviceSynchronize();
  free_buffers();
  cudaFree(d_gptr_); cudaFree(d_gval_); cudaFree(d_gcol_); cudaFree(d_diag4_); cudaFree(d_rr4_);
  cudaFree(d_vec_); cudaFree(d_mask_); cudaFree(d_map_); cudaFree(d_total_); cudaFree(d_rec_); cudaFree(d_stage_);
  cudaFreeHost(h_total_);
  if (ev_total_) cudaEventDestroy((cudaEvent_t)ev_total_);
  for (auto& e : ev_pending_) { cudaEventDestroy((cudaEvent_t)e.start); cudaEventDestroy((cudaEvent_t)e.stop); }
  for (void* e : ev_pool_) cudaEventDestroy((cudaEvent_t)e);
  if (own_stream_ && stream_) cudaStreamDestroy(ST);
}

void Engine::set_stream(void* s) {
  CK(cudaSetDevice(device_));
  if (stream_) CK(cudaStreamSynchronize(ST));
  if (own_stream_ && stream_) cudaStreamDestroy(ST);
  stream_ = s;
  own_stream_ = false;
}

void Engine::sync() {
  CK(cudaSetDevice(device_));
  CK(cudaStreamSynchronize(ST));
}

void Engine::free_buffers() {
  for (int i = 0; i < 2; ++i) {
    if (buf_[i]) cudaFree(buf_[i]);
    buf_[i] = nullptr;
  }
}

size_t Engine::bytes_allocated() const {
  const size_t per = (size_t)(ny_ + 2) * row_stride() * sizeof(int32_t);
  return (buf_[0] ? per : 0) + (buf_[1] ? per : 0);
}

int32_t* Engine::row0(int which) const { return buf_[which] + row_stride(); }

void Engine::alloc_buffers(uint32_t cap) {
  free_buffers();
  cap_ = cap;
  cur_ = 0;
  if (cap_ == 0) return;
  const size_t bytes = (size_t)(ny_ + 2) * row_stride() * sizeof(int32_t);
  CK(cudaMalloc(&buf_[0], bytes));
  CK(cudaMemsetAsync(buf_[0], 0, bytes, ST));
}

void Engine::ensure_other() {
  if (buf_[cur_ ^ 1]) return;
  const size_t bytes = (size_t)(ny_ + 2) * row_stride() * sizeof(int32_t);
  CK(cudaMalloc(&buf_[cur_ ^ 1], bytes));
  CK(cudaMemsetAsync(buf_[cur_ ^ 1], 0, bytes, ST));
}

void Engine::set_grid(uint32_t nx, uint32_t ny_global, uint32_t y0, uint32_t ny, uint32_t nsp) {
  CK(cudaSetDevice(device_));
  sync();
  nx_ = nx; nyg_ = ny_global; y0_ = y0; ny_ = ny;
  ntx_ = (nx + 31) / 32;
  nsp_ = nsp;
  peer_top_ = peer_bottom_ = nullptr;
  alloc_buffers(round_up4(nsp));
}

void Engine::resize_species(uint32_t nsp) {
  CK(cudaSetDevice(device_));
  if (nsp <= cap_) {
    if (nsp > nsp_ && buf_[cur_]) {
      const uint32_t ntiles = ntx_ * ny_;
      zero_rows_kernel<<<std::min<unsigned long long>(4096, (unsigned long long)ntiles * (nsp - nsp_) / 8 + 1), 256, 0, ST>>>(
          row0(cur_), tile_stride(), ntiles, nsp_, nsp);
      CK(cudaGetLastError());
      ++launches_;
    }
    nsp_ = nsp;
    return;
  }
  // grow: new capacity with head-room, re-layout into a fresh buffer
  const uint32_t new_cap = round_up4(nsp + nsp / 4 + 16);
  const size_t old_stride = tile_stride();
  int32_t* old_cur = buf_[cur_];
  int32_t* old_other = buf_[cur_ ^ 1];
  if (old_other) { CK(cudaStreamSynchronize(ST)); CK(cudaFree(old_other)); }
  buf_[0] = buf_[1] = nullptr;
  const uint32_t ntiles_all = ntx_ * (ny_ + 2);
  const uint32_t old_nsp = nsp_;
  cap_ = new_cap;
  const size_t bytes = (size_t)(ny_ + 2) * row_stride() * sizeof(int32_t);
  int32_t* nb = nullptr;
  CK(cudaMalloc(&nb, bytes));
  if (old_cur) {
    relayout_kernel<<<std::min<unsigned long long>(8192, (unsigned long long)ntiles_all * new_cap / 8 + 1), 256, 0, ST>>>(
        old_cur, old_stride, nb, tile_stride(), ntiles_all, old_nsp, new_cap);
    CK(cudaGetLastError());
    ++launches_;
    CK(cudaStreamSynchronize(ST));
    CK(cudaFree(old_cur));
  } else {
    CK(cudaMemsetAsync(nb, 0, bytes, ST));
  }
  buf_[0] = nb;
  cur_ = 0;
  nsp_ = nsp;
  peer_top_ = peer_bottom_ = nullptr;
}

// --- density access ----------------------------------------------------------
void Engine::set_density(size_t cell, const int32_t* d, size_t n) {
  CK(cudaSetDevice(device_));
  if (n > nsp_) throw std::runtime_error("set_density: more species than the model has");
  if (cell >= ncells_local()) throw std::runtime_error("set_density: cell out of range");
  const uint32_t yl = (uint32_t)(cell / nx_), x = (uint32_t)(cell % nx_);
  int32_t* dst = row0(cur_) + ((size_t)yl * ntx_ + x / 32) * tile_stride() + (x & 31u);
  CK(cudaMemcpy2DAsync(dst, 128, d, 4, 4, n, cudaMemcpyHostToDevice, ST));
  if (n < nsp_) CK(cudaMemset2DAsync(dst + n * 32, 128, 0, 4, nsp_ - n, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::get_density(size_t cell, int32_t* d, size_t n) {
  CK(cudaSetDevice(device_));
  if (n > nsp_) throw std::runtime_error("get_density: more species than the model has");
  if (cell >= ncells_local()) throw std::runtime_error("get_density: cell out of range");
  const uint32_t yl = (uint32_t)(cell / nx_), x = (uint32_t)(cell % nx_);
  const int32_t* src = row0(cur_) + ((size_t)yl * ntx_ + x / 32) * tile_stride() + (x & 31u);
  CK(cudaMemcpy2DAsync(d, 4, src, 128, 4, n, cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::set_density_all_dev(const int32_t* dev, size_t nsp) {
  CK(cudaSetDevice(device_));
  if (nsp != nsp_) throw std::runtime_error("set_density_all: species count mismatch");
  const unsigned long long nwork = (unsigned long long)ntx_ * ny_ * ((nsp_ + 31) / 32);
  if (!nwork) return;
  transpose_kernel<true><<<(unsigned)std::min<unsigned long long>(nwork, 148ull * 64), 1024, 0, ST>>>(
      row0(cur_), tile_stride(), ntx_, nx_, ny_, nsp_, const_cast<int32_t*>(dev));
  CK(cudaGetLastError());
  ++launches_;
}

void Engine::get_density_all_dev(int32_t* dev, size_t nsp) {
  CK(cudaSetDevice(device_));
  if (nsp != nsp_) throw std::runtime_error("get_density_all: species count mismatch");
  const unsigned long long nwork = (unsigned long long)ntx_ * ny_ * ((nsp_ + 31) / 32);
  if (!nwork) return;
  transpose_kernel<false><<<(unsigned)std::min<unsigned long long>(nwork, 148ull * 64), 1024, 0, ST>>>(
      row0(cur_), tile_stride(), ntx_, nx_, ny_, nsp_, dev);
  CK(cudaGetLastError());
  ++launches_;
}

void Engine::set_density_all(const int32_t* host, size_t nsp) {
  CK(cudaSetDevice(device_));
  const size_t n = ncells_local() * nsp;
  if (n > stage_cap_) {
    if (d_stage_) CK(cudaFree(d_stage_));
    CK(cudaMalloc(&d_stage_, n * sizeof(int32_t)));
    stage_cap_ = n;
  }
  CK(cudaMemcpyAsync(d_stage_, host, n * sizeof(int32_t), cudaMemcpyHostToDevice, ST));
  set_density_all_dev(d_stage_, nsp);
  CK(cudaStreamSynchronize(ST));
}

void Engine::get_density_all(int32_t* host, size_t nsp) {
  CK(cudaSetDevice(device_));
  const size_t n = ncells_local() * nsp;
  if (n > stage_cap_) {
    if (d_stage_) CK(cudaFree(d_stage_));
    CK(cudaMalloc(&d_stage_, n * sizeof(int32_t)));
    stage_cap_ = n;
  }
  get_density_all_dev(d_stage_, nsp);
  CK(cudaMemcpyAsync(host, d_stage_, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::fill_density(int32_t v) {
  CK(cudaSetDevice(device_));
  const uint32_t ntiles = ntx_ * ny_;
  if (!ntiles || !nsp_) return;
  fill_kernel<<<std::min<unsigned long long>(8192, (unsigned long long)ntiles * nsp_ / 8 + 1), 256, 0, ST>>>(
      row0(cur_), tile_stride(), ntiles, ntx_, nx_, nsp_, v);
  CK(cudaGetLastError());
  ++launches_;
}

void Engine::fill_density_random(uint32_t seed, int32_t lo, int32_t hi) {
  CK(cudaSetDevice(device_));
  const uint32_t ntiles = ntx_ * ny_;
  if (!ntiles || !nsp_) return;
  const unsigned long long nwork = (unsigned long long)ntiles * ((nsp_ + 3) / 4);
  fill_random_kernel<<<(unsigned)std::min<unsigned long long>((nwork + 7) / 8, 148ull * 32), 256, 0, ST>>>(
      row0(cur_), tile_stride(), ntiles, ntx_, nx_, y0_, nsp_, seed, lo, (uint32_t)(hi - lo) + 1u);
  CK(cudaGetLastError());
  ++launches_;
}

// --- beta ----------------------------------------------------------------------
void Engine::upload_beta(const BetaGell& b) {
  CK(cudaSetDevice(device_));
  const size_t nslots = b.gptr.empty() ? 0 : b.gptr.back();
  if (b.ngroups + 1 > beta_cap_groups_) {
    CK(cudaStreamSynchronize(ST));
    cudaFree(d_gptr_); cudaFree(d_diag4_); cudaFree(d_rr4_);
    beta_cap_groups_ = (size_t)(b.ngroups + 1) * 5 / 4 + 16;
    CK(cudaMalloc(&d_gptr_, beta_cap_groups_ * sizeof(uint32_t)));
    CK(cudaMalloc(&d_diag4_, beta_cap_groups_ * 4 * sizeof(double)));
    CK(cudaMalloc(&d_rr4_, beta_cap_groups_ * 4 * sizeof(double)));
  }
  if (nslots + 1 > beta_cap_slots_) {
    CK(cudaStreamSynchronize(ST));
    cudaFree(d_gval_); cudaFree(d_gcol_);
    beta_cap_slots_ = (nslots + 1) * 5 / 4 + 64;
    CK(cudaMalloc(&d_gval_, beta_cap_slots_ * 4 * sizeof(double)));
    CK(cudaMalloc(&d_gcol_, beta_cap_slots_ * 4 * sizeof(uint32_t)));
  }
  // synchronous copies: the host vectors are caller-owned temporaries
  CK(cudaStreamSynchronize(ST));
  CK(cudaMemcpy(d_gptr_, b.gptr.data(), (b.ngroups + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice));
  if (b.ngroups) {
    CK(cudaMemcpy(d_diag4_, b.diag4.data(), (size_t)b.ngroups * 4 * sizeof(double), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_rr4_, b.rr4.data(), (size_t)b.ngroups * 4 * sizeof(double), cudaMemcpyHostToDevice));
  }
  if (nslots) {
    CK(cudaMemcpy(d_gval_, b.gval.data(), nslots * 4 * sizeof(double), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_gcol_, b.gcol.data(), nslots * 4 * sizeof(uint32_t), cudaMemcpyHostToDevice));
  }
  ngroups_ = b.ngroups;
  max_width_ = 0;
  for (uint32_t g = 0; g < b.ngroups; ++g) max_width_ = std::max(max_width_, b.gptr[g + 1] - b.gptr[g]);
}

// --- generate --------------------------------------------------------------------
template <int THREADS, bool DB>
static void launch_generate(const GenParams& p, int grid, size_t smem, cudaStream_t st) {
  CK(cudaFuncSetAttribute(generate_kernel<THREADS, DB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  generate_kernel<THREADS, DB><<<grid, THREADS, smem, st>>>(p);
}

// ---- optional per-kernel timing (CUDA events on the engine's stream) ----------------
void Engine::time_begin(int kind) {
  if (!timing) return;
  cudaEvent_t a, b;
  if (ev_pool_.size() >= 2) {
    a = (cudaEvent_t)ev_pool_.back(); ev_pool_.pop_back();
    b = (cudaEvent_t)ev_pool_.back(); ev_pool_.pop_back();
  } else {
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
  }
  CK(cudaEventRecord(a, ST));
  ev_pending_.push_back(PendingEvent{a, b, kind});
}
void Engine::time_end() {
  if (!timing) return;
  CK(cudaEventRecord((cudaEvent_t)ev_pending_.back().stop, ST));
}
void Engine::kernel_times(double* ms, uint64_t* count) {
  CK(cudaSetDevice(device_));
  CK(cudaStreamSynchronize(ST));
  for (auto& e : ev_pending_) {
    float t = 0;
    CK(cudaEventElapsedTime(&t, (cudaEvent_t)e.start, (cudaEvent_t)e.stop));
    kt_ms_[e.kind] += t;
    kt_n_[e.kind] += 1;
    ev_pool_.push_back(e.start);
    ev_pool_.push_back(e.stop);
  }
  ev_pending_.clear();
  for (int k = 0; k < K_COUNT; ++k) { ms[k] = kt_ms_[k]; count[k] = kt_n_[k]; }
}
void Engine::reset_kernel_times() {
  double ms[K_COUNT]; uint64_t n[K_COUNT];
  kernel_times(ms, n);
  for (int k = 0; k < K_COUNT; ++k) { kt_ms_[k] = 0; kt_n_[k] = 0; }
}

static void fill_gen_params(GenParams& p, int32_t* dens, size_t tile_stride, uint32_t ntx, uint32_t ny, uint32_t nx,
                            uint32_t y0, uint32_t nsp, uint32_t ngroups, const uint32_t* gptr, const double* gval,
                            const uint32_t* gcol, const double* diag4, const double* rr4, uint32_t seed) {
  p.dens = dens;
  p.tile_stride = tile_stride;
  p.ntiles = ntx * ny;
  p.ntx = ntx; p.nx = nx; p.y0 = y0;
  p.nsp = nsp; p.ngroups = ngroups;
  p.gptr = gptr; p.gval = gval; p.gcol = gcol; p.diag4 = diag4; p.rr4 = rr4;
  p.seed = seed;
  p.niter = 1;
  p.t0 = 0;
}

void Engine::pre_rounding(double* out_host) {
  CK(cudaSetDevice(device_));
  if (!nsp_ || !buf_[cur_]) return;
  if (ngroups_ != (nsp_ + 3) / 4) throw std::runtime_error("pre_rounding: beta not uploaded for the current species count");
  GenParams p;
  fill_gen_params(p, row0(cur_), tile_stride(), ntx_, ny_, nx_, y0_, nsp_, ngroups_, d_gptr_, d_gval_, d_gcol_, d_diag4_,
                  d_rr4_, 0);
  const size_t n = ncells_local() * nsp_;
  double* d_out = nullptr;
  CK(cudaMalloc(&d_out, n * sizeof(double)));
  const unsigned long long nwork = (unsigned long long)p.ntiles * p.ngroups;
  pre_rounding_kernel<<<(unsigned)std::min<unsigned long long>((nwork + 7) / 8, 148ull * 16), 256, 0, ST>>>(p, d_out);
  CK(cudaGetLastError());
  ++launches_;
  CK(cudaMemcpyAsync(out_host, d_out, n * sizeof(double), cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
  CK(cudaFree(d_out));
}

void Engine::generate(uint32_t niter, uint64_t t0, uint32_t seed) {
  CK(cudaSetDevice(device_));
  if (!niter || !nsp_ || !buf_[cur_]) return;
  if (ngroups_ != (nsp_ + 3) / 4) throw std::runtime_error("generate: beta not uploaded for the current species count");
  GenParams p;
  fill_gen_params(p, row0(cur_), tile_stride(), ntx_, ny_, nx_, y0_, nsp_, ngroups_, d_gptr_, d_gval_, d_gcol_, d_diag4_,
                  d_rr4_, seed);
  const size_t tile_bytes = (size_t)ngroups_ * 4 * 32 * sizeof(int32_t);
  const size_t smem1 = tile_bytes + 16, smem2 = 2 * tile_bytes + 16;
  if (smem1 > (size_t)max_smem_)
    throw std::runtime_error("generate: species count exceeds the shared-memory tile (nsp*128 B must fit in 227 KB)");
  const bool can_db = smem2 <= (size_t)max_smem_;
  const int threads = gen_variant == 256 || gen_variant == 1024 ? gen_variant : 512;
  auto launch = [&](bool db, uint32_t n_it, uint64_t t) {
    p.niter = n_it;
    p.t0 = t;
    const size_t smem = db ? smem2 : smem1;
    // resident CTAs per SM limited by shared memory (and 2048 threads/SM)
    int per_sm = (int)std::max<size_t>(1, std::min<size_t>((size_t)(227 * 1024) / (smem + 1024), (size_t)(2048 / threads)));
    const int grid = (int)std::min<unsigned long long>(p.ntiles, (unsigned long long)num_sms_ * per_sm);
    time_begin(K_GENERATE);
    if (threads == 256) db ? launch_generate<256, true>(p, grid, smem, ST) : launch_generate<256, false>(p, grid, smem, ST);
    else if (threads == 1024) db ? launch_generate<1024, true>(p, grid, smem, ST) : launch_generate<1024, false>(p, grid, smem, ST);
    else db ? launch_generate<512, true>(p, grid, smem, ST) : launch_generate<512, false>(p, grid, smem, ST);
    time_end();
    CK(cudaGetLastError());
    ++launches_;
    last_grid_ = grid; last_block_ = threads; last_smem_ = (int)smem;
  };
  if (niter == 1) launch(false, 1, t0);
  else if (can_db) launch(true, niter, t0);
  else for (uint32_t i = 0; i < niter; ++i) launch(false, 1, t0 + i);
}

// --- migrate ---------------------------------------------------------------------
uint64_t Engine::migrate(const double* cm_host, bool wait_result) {
  CK(cudaSetDevice(device_));
  if (!nsp_ || !buf_[cur_]) return 0;
  ensure_other();
  if (nsp_ > vec_cap_) {
    CK(cudaStreamSynchronize(ST));
    cudaFree(d_vec_);
    vec_cap_ = (size_t)nsp_ * 5 / 4 + 64;
    CK(cudaMalloc(&d_vec_, vec_cap_ * sizeof(double)));
  }
  CK(cudaMemcpyAsync(d_vec_, cm_host, nsp_ * sizeof(double), cudaMemcpyHostToDevice, ST));
  CK(cudaMemsetAsync(d_total_, 0, sizeof(unsigned long long), ST));
  MigParams p;
  p.src = row0(cur_);
  p.dst = row0(cur_ ^ 1);
  p.top_row = peer_top_ ? (const int32_t*)peer_top_ : buf_[cur_];
  p.bottom_row = peer_bottom_ ? (const int32_t*)peer_bottom_ : row0(cur_) + (size_t)ny_ * row_stride();
  p.tile_stride = tile_stride(); p.row_stride = row_stride();
  p.ntx = ntx_; p.nx = nx_; p.ny = ny_; p.y0 = y0_; p.nyg = nyg_; p.nsp = nsp_;
  p.chunk = 64;
  p.nchunks = (nsp_ + p.chunk - 1) / p.chunk;
  p.cm = d_vec_;
  p.total = d_total_;
  const unsigned long long nwork = (unsigned long long)ntx_ * ny_ * p.nchunks;
  const int grid = (int)std::min<unsigned long long>((nwork + 7) / 8, (unsigned long long)num_sms_ * 8);
  time_begin(K_MIGRATE);
  migrate_kernel<<<grid, 256, 0, ST>>>(p);
  time_end();
  CK(cudaGetLastError());
  ++launches_;
  last_grid_ = grid; last_block_ = 256; last_smem_ = 0;
  cur_ ^= 1;
  CK(cudaMemcpyAsync(h_total_, d_total_, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ST));
  CK(cudaEventRecord((cudaEvent_t)ev_total_, ST));
  return wait_result ? migrate_result() : 0;
}

uint64_t Engine::migrate_result() {
  CK(cudaEventSynchronize((cudaEvent_t)ev_total_));
  return h_total_[0];
}

// --- condense ----------------------------------------------------------------------
void Engine::alive_mask(uint8_t* alive_host) {
  CK(cudaSetDevice(device_));
  if (!nsp_) return;
  if (nsp_ > mask_cap_) {
    CK(cudaStreamSynchronize(ST));
    cudaFree(d_mask_); cudaFree(d_map_);
    mask_cap_ = (size_t)nsp_ * 5 / 4 + 64;
    CK(cudaMalloc(&d_mask_, mask_cap_));
    CK(cudaMalloc(&d_map_, mask_cap_ * sizeof(uint32_t)));
  }
  CK(cudaMemsetAsync(d_mask_, 0, nsp_, ST));
  const uint32_t ntiles = ntx_ * ny_;
  if (ntiles) {
    alive_kernel<<<std::min<uint32_t>(ntiles, num_sms_ * 8), 256, 0, ST>>>(row0(cur_), tile_stride(), ntiles, nsp_, d_mask_);
    CK(cudaGetLastError());
    ++launches_;
  }
  CK(cudaMemcpyAsync(alive_host, d_mask_, nsp_, cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::compact_species(const uint8_t* keep_host, uint32_t nkeep) {
  CK(cudaSetDevice(device_));
  if (!nsp_) return;
  std::vector<uint32_t> map(nsp_);
  uint32_t k = 0;
  for (uint32_t s = 0; s < nsp_; ++s) { map[s] = k; if (keep_host[s]) ++k; }
  if (k != nkeep) throw std::runtime_error("compact_species: mask/count mismatch");
  if (nsp_ > mask_cap_) {
    CK(cudaStreamSynchronize(ST));
    cudaFree(d_mask_); cudaFree(d_map_);
    mask_cap_ = (size_t)nsp_ * 5 / 4 + 64;
    CK(cudaMalloc(&d_mask_, mask_cap_));
    CK(cudaMalloc(&d_map_, mask_cap_ * sizeof(uint32_t)));
  }
  CK(cudaMemcpyAsync(d_mask_, keep_host, nsp_, cudaMemcpyHostToDevice, ST));
  CK(cudaMemcpyAsync(d_map_, map.data(), nsp_ * sizeof(uint32_t), cudaMemcpyHostToDevice, ST));
  const uint32_t ntiles_all = ntx_ * (ny_ + 2);   // ghost rows too: keeps halos consistent
  compact_kernel<<<std::min<uint32_t>(ntiles_all, num_sms_ * 2), 1024, 0, ST>>>(buf_[cur_], tile_stride(), ntiles_all,
                                                                                nsp_, nkeep, d_mask_, d_map_);
  CK(cudaGetLastError());
  ++launches_;
  CK(cudaStreamSynchronize(ST));   // map is a host temporary
  nsp_ = nkeep;
}

// --- mutate ------------------------------------------------------------------------
void Engine::mutate_cells(const double* ms_host, uint64_t t, uint32_t seed, std::vector<MutRecord>& out) {
  CK(cudaSetDevice(device_));
  out.clear();
  if (!nsp_ || !buf_[cur_]) return;
  if (nsp_ > vec_cap_) {
    CK(cudaStreamSynchronize(ST));
    cudaFree(d_vec_);
    vec_cap_ = (size_t)nsp_ * 5 / 4 + 64;
    CK(cudaMalloc(&d_vec_, vec_cap_ * sizeof(double)));
  }
  CK(cudaMemcpyAsync(d_vec_, ms_host, nsp_ * sizeof(double), cudaMemcpyHostToDevice, ST));
  if (!d_rec_) {
    rec_cap_ = 1u << 16;
    CK(cudaMalloc(&d_rec_, rec_cap_ * sizeof(MutRecord)));
  }
  const uint32_t ntiles = ntx_ * ny_;
  unsigned long long count = 0;
  for (int attempt = 0; attempt < 2; ++attempt) {
    CK(cudaMemsetAsync(d_total_ + 1, 0, sizeof(unsigned long long), ST));
    mutate_scan_kernel<<<std::min<uint32_t>(ntiles, num_sms_ * 8), 256, 0, ST>>>(
        row0(cur_), tile_stride(), ntiles, ntx_, nx_, y0_, nsp_, d_vec_, seed, t, d_rec_, rec_cap_, d_total_ + 1);
    CK(cudaGetLastError());
    ++launches_;
    CK(cudaMemcpyAsync(h_total_ + 1, d_total_ + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ST));
    CK(cudaStreamSynchronize(ST));
    count = h_total_[1];
    if (count <= rec_cap_) break;
    if (count > 0xfffffff0ull) throw std::runtime_error("mutate: more than 2^32 speciation events in one call");
    cudaFree(d_rec_);
    rec_cap_ = count + count / 8 + 1024;
    CK(cudaMalloc(&d_rec_, rec_cap_ * sizeof(MutRecord)));
  }
  if (!count) return;
  mutate_apply_kernel<<<(unsigned)((count + 255) / 256), 256, 0, ST>>>(row0(cur_), tile_stride(), ntx_, nx_, y0_, d_rec_,
                                                                       (uint32_t)count);
  CK(cudaGetLastError());
  ++launches_;
  out.resize(count);
  CK(cudaMemcpyAsync(out.data(), d_rec_, count * sizeof(MutRecord), cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
  std::sort(out.begin(), out.end(), [](const MutRecord& a, const MutRecord& b) {
    return a.cell != b.cell ? a.cell < b.cell : a.slot < b.slot;
  });
}

void Engine::append_species(const uint32_t* cells, size_t m) {
  CK(cudaSetDevice(device_));
  if (!m) return;
  const uint32_t s0 = nsp_;
  resize_species(nsp_ + (uint32_t)m);   // zero-fills the new rows
  if (!buf_[cur_]) return;
  uint32_t* d_cells = nullptr;
  CK(cudaMalloc(&d_cells, m * sizeof(uint32_t)));
  CK(cudaMemcpyAsync(d_cells, cells, m * sizeof(uint32_t), cudaMemcpyHostToDevice, ST));
  birth_kernel<<<(unsigned)((m + 255) / 256), 256, 0, ST>>>(row0(cur_), tile_stride(), ntx_, nx_, y0_, ny_, s0, d_cells,
                                                            (uint32_t)m);
  CK(cudaGetLastError());
  ++launches_;
  CK(cudaStreamSynchronize(ST));
  CK(cudaFree(d_cells));
}

// --- observables ------------------------------------------------------------------
void Engine::nsp_per_cell(uint32_t* out_host) {
  CK(cudaSetDevice(device_));
  const size_t n = ncells_local();
  if (!n) return;
  if (n > stage_cap_) {
    if (d_stage_) CK(cudaFree(d_stage_));
    CK(cudaMalloc(&d_stage_, n * sizeof(int32_t)));
    stage_cap_ = n;
  }
  const uint32_t ntiles = ntx_ * ny_;
  if (nsp_ && buf_[cur_]) {
    nsp_cell_kernel<<<std::min<uint32_t>(ntiles, num_sms_ * 8), 256, 0, ST>>>(row0(cur_), tile_stride(), ntiles, ntx_, nx_,
                                                                              nsp_, (uint32_t*)d_stage_);
    CK(cudaGetLastError());
    ++launches_;
  } else {
    CK(cudaMemsetAsync(d_stage_, 0, n * sizeof(uint32_t), ST));
  }
  CK(cudaMemcpyAsync(out_host, d_stage_, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::species_totals(int64_t* out_host) {
  CK(cudaSetDevice(device_));
  if (!nsp_) return;
  long long* d = nullptr;
  CK(cudaMalloc(&d, nsp_ * sizeof(long long)));
  totals_kernel<<<std::min<uint32_t>((nsp_ + 7) / 8, 4096), 256, 0, ST>>>(row0(cur_), tile_stride(), ntx_ * ny_, nsp_, d);
  CK(cudaGetLastError());
  ++launches_;
  CK(cudaMemcpyAsync(out_host, d, nsp_ * sizeof(long long), cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
  CK(cudaFree(d));
}

// --- halo ---------------------------------------------------------------------------
void Engine::halo_ptrs(void** send_top, void** send_bottom, void** recv_top, void** recv_bottom) {
  int32_t* r0 = row0(cur_);
  *send_top = r0;
  *send_bottom = r0 + (size_t)(ny_ - 1) * row_stride();
  *recv_top = buf_[cur_];
  *recv_bottom = r0 + (size_t)ny_ * row_stride();
}

void Engine::set_peer_rows(const void* top, const void* bottom) {
  peer_top_ = top;
  peer_bottom_ = bottom;
}

int Engine::ipc_export(void* handle_out64, size_t* offset_top, size_t* offset_bottom) {
  CK(cudaSetDevice(device_));
  cudaIpcMemHandle_t h;
  CK(cudaIpcGetMemHandle(&h, buf_[cur_]));
  static_assert(sizeof(h) == 64, "ipc handle size");
  memcpy(handle_out64, &h, 64);
  *offset_top = row_stride() * sizeof(int32_t);                       // first owned row
  *offset_bottom = (size_t)ny_ * row_stride() * sizeof(int32_t);      // last owned row
  return 0;
}

void* Engine::ipc_import(const void* handle64) {
  CK(cudaSetDevice(device_));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  return p;
}

}  // namespace eco
