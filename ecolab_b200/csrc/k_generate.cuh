// k_generate.cuh -- the generalised Lotka-Volterra step with stochastic rounding
// (reference EcolabPoint::generate, models/ecolab_model.cc:90-112; ROUND :57-66).
//
// beta on the device is a blocked ELL matrix matched to the lane map: species rows are processed
// ROWS = 128/TW at a time (one warp-op); block b has width = longest row in the block, shorter rows
// padded with (val 0, offset 0).  Entry (slot j, row r) is the 16-byte record
//     { double val ; uint32 byte offset of species row `col` inside a tile ; pad }
// at ell[j*ROWS + r], so the lanes of one row fetch one broadcast 16-byte load per term.  Terms of a
// row keep the COO order (reference oDiagIdx, :718-723) so the FP64 summation order is the reference's:
//     ir = diag*n ; ir += val_j * n[col_j] (j ascending) ; x = n + n*(r + ir), every op individually rounded.
#pragma once
#include "k_common.cuh"

namespace eco {

struct GenParams {
  Layout L;
  uint32_t nblocks;          // ceil(nsp / ROWS)
  const uint32_t* bptr;      // [nblocks+1] slot offsets
  const uint4* ell;          // [nslots*ROWS]
  const double2* rowpar;     // [nblocks*ROWS] {diag, repro_rate}
  uint32_t seed, niter;
  unsigned long long t0;
};

// One warp-op: ROWS species rows x TW cells; this lane owns 4 consecutive cells of one row.
// `lp` = shared-memory tile + 4*quad (ints); returns the four pre-rounding values.
template <int TW>
__device__ __forceinline__ void gen_x(const int32_t* __restrict__ lp, uint32_t row, uint32_t rr, uint32_t b,
                                      const GenParams& p, double x[4]) {
  constexpr int R = LaneMap<TW>::ROWS;
  const int4 own = *reinterpret_cast<const int4*>(lp + row * TW);
  const double d0 = u2d((uint32_t)own.x), d1 = u2d((uint32_t)own.y), d2 = u2d((uint32_t)own.z), d3 = u2d((uint32_t)own.w);
  const double2 par = __ldg(p.rowpar + row);
  double ir0 = __dmul_rn(par.x, d0), ir1 = __dmul_rn(par.x, d1), ir2 = __dmul_rn(par.x, d2), ir3 = __dmul_rn(par.x, d3);
  uint32_t j = __ldg(p.bptr + b);
  const uint32_t je = __ldg(p.bptr + b + 1);
  const uint4* e = p.ell + (size_t)j * R + rr;
  for (; j < je; ++j, e += R) {
    const uint4 w = __ldg(e);
    const double v = __hiloint2double((int)w.y, (int)w.x);
    const int4 g = *reinterpret_cast<const int4*>(reinterpret_cast<const char*>(lp) + w.z);
    ir0 = __dadd_rn(ir0, __dmul_rn(v, u2d((uint32_t)g.x)));
    ir1 = __dadd_rn(ir1, __dmul_rn(v, u2d((uint32_t)g.y)));
    ir2 = __dadd_rn(ir2, __dmul_rn(v, u2d((uint32_t)g.z)));
    ir3 = __dadd_rn(ir3, __dmul_rn(v, u2d((uint32_t)g.w)));
  }
  x[0] = __dadd_rn(d0, __dmul_rn(d0, __dadd_rn(par.y, ir0)));
  x[1] = __dadd_rn(d1, __dmul_rn(d1, __dadd_rn(par.y, ir1)));
  x[2] = __dadd_rn(d2, __dmul_rn(d2, __dadd_rn(par.y, ir2)));
  x[3] = __dadd_rn(d3, __dmul_rn(d3, __dadd_rn(par.y, ir3)));
}

// Persistent CTAs; each iteration: bulk-copy one tile (all species of TW cells) into shared memory, run
// `niter` Jacobi steps, write the tile back in place.
//  DB=false: niter==1, one smem buffer, results stored straight to HBM as 16-byte vectors (a warp-op
//            writes 512 contiguous bytes); several CTAs are resident per SM so one CTA's bulk load
//            overlaps the others' arithmetic.
//  DB=true : two smem buffers, any niter (HBM traffic amortised over niter).
template <int TW, int THREADS, int MINB, bool DB>
__global__ void __launch_bounds__(THREADS, MINB) generate_kernel(const GenParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int R = LaneMap<TW>::ROWS;
  const uint32_t rows = p.nblocks * R;
  const uint32_t tile_words = rows * TW;
  const uint32_t tile_bytes = tile_words * 4u;
  int32_t* buf0 = reinterpret_cast<int32_t*>(smem_raw);
  int32_t* buf1 = buf0 + (DB ? tile_words : 0);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + (size_t)tile_bytes * (DB ? 2 : 1));

  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = THREADS / 32;
  const uint32_t rr = LaneMap<TW>::row(lane), q = LaneMap<TW>::quad(lane);
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const uint32_t ntiles = p.L.ntx * p.L.ny;
  uint32_t phase = 0;
  for (uint32_t tile_id = blockIdx.x; tile_id < ntiles; tile_id += gridDim.x) {
    int32_t* gt = p.L.base + (size_t)tile_id * p.L.tile_stride;
    if (threadIdx.x == 0) {
      mbar_expect_tx(bar, tile_bytes);
      for (uint32_t off = 0; off < tile_bytes; off += 32768u) {   // modest requests, one barrier
        const uint32_t n = min(32768u, tile_bytes - off);
        bulk_g2s(reinterpret_cast<unsigned char*>(buf0) + off, reinterpret_cast<const unsigned char*>(gt) + off, n, bar);
      }
      const uint32_t nxt = tile_id + gridDim.x;
      if (nxt < ntiles) bulk_prefetch_l2(p.L.base + (size_t)nxt * p.L.tile_stride, tile_bytes);
    }
    mbar_wait(bar, phase);
    phase ^= 1u;

    const uint32_t yl = tile_id / p.L.ntx, tx = tile_id - yl * p.L.ntx;
    const uint32_t xq = tx * (TW / 4) + q, yg = p.L.y0 + yl;

    if (!DB) {
      const int32_t* lp = buf0 + 4 * q;
      for (uint32_t b = warp; b < p.nblocks; b += nwarps) {
        const uint32_t row = b * R + rr;
        double x[4];
        gen_x<TW>(lp, row, rr, b, p, x);
        const uint4 u = draw4(xq, yg, row, p.t0, p.seed, 0u);
        int4 o;
        o.x = round_stoch(x[0], u.x); o.y = round_stoch(x[1], u.y);
        o.z = round_stoch(x[2], u.z); o.w = round_stoch(x[3], u.w);
        *reinterpret_cast<int4*>(gt + row * TW + 4 * q) = o;
      }
      __syncthreads();   // everyone done reading buf0 before the next bulk copy lands
    } else {
      int32_t* src = buf0;
      int32_t* dst = buf1;
      for (uint32_t step = 0; step < p.niter; ++step) {
        const int32_t* lp = src + 4 * q;
        for (uint32_t b = warp; b < p.nblocks; b += nwarps) {
          const uint32_t row = b * R + rr;
          double x[4];
          gen_x<TW>(lp, row, rr, b, p, x);
          const uint4 u = draw4(xq, yg, row, p.t0 + step, p.seed, 0u);
          int4 o;
          o.x = round_stoch(x[0], u.x); o.y = round_stoch(x[1], u.y);
          o.z = round_stoch(x[2], u.z); o.w = round_stoch(x[3], u.w);
          *reinterpret_cast<int4*>(dst + row * TW + 4 * q) = o;
        }
        __syncthreads();
        int32_t* tmp = src; src = dst; dst = tmp;
      }
      const int4* s4 = reinterpret_cast<const int4*>(src);
      int4* g4 = reinterpret_cast<int4*>(gt);
      for (uint32_t i = threadIdx.x; i < tile_words / 4; i += THREADS) g4[i] = s4[i];
      __syncthreads();
    }
  }
}

// x = n + n*(r + ir) of ONE generate sub-step, before ROUND (reference :99-102), for every owned cell;
// densities are not modified.  Diagnostic path for the "pre-rounding float update within 1e-6"
// criterion: one thread per (cell, species), operands straight from global memory, runtime tile width.
__global__ void __launch_bounds__(256) pre_rounding_kernel(const GenParams p, uint32_t R, double* __restrict__ out) {
  const Layout& L = p.L;
  const unsigned long long n = (unsigned long long)L.nx * L.ny * L.nsp;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t s = (uint32_t)(i % L.nsp);
    const unsigned long long c = i / L.nsp;
    const uint32_t x = (uint32_t)(c % L.nx), yl = (uint32_t)(c / L.nx);
    const int32_t* tile = L.base + ((size_t)yl * L.ntx + x / L.tw) * L.tile_stride + x % L.tw;
    const double d = i2d(tile[(size_t)s * L.tw]);
    const double2 par = p.rowpar[s];
    double ir = __dmul_rn(par.x, d);
    const uint32_t b = s / R, rr = s % R;
    for (uint32_t j = p.bptr[b]; j < p.bptr[b + 1]; ++j) {
      const uint4 w = p.ell[(size_t)j * R + rr];
      const double v = __hiloint2double((int)w.y, (int)w.x);
      ir = __dadd_rn(ir, __dmul_rn(v, i2d(tile[w.z / 4])));   // w.z = col*TW*4 bytes
    }
    out[i] = __dadd_rn(d, __dmul_rn(d, __dadd_rn(par.y, ir)));
  }
}

}  // namespace eco
