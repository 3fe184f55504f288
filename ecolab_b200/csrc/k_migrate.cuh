// k_migrate.cuh -- nearest-neighbour migration (reference SpatialModel::migrate,
// models/ecolab_model.cc:616-703):
//     delta(c,s) = sum over existing W,E,N,S neighbours b of (int)(cm[s] * (double)(n_b - n_c))   (:631-642)
//     n_new = n_old + delta  (synchronous: reads `src`, writes `dst`)                              (:649-664)
//     returns sum |delta|                                                                          (:660, :702)
// The pairwise term T(a,b) = trunc(cm*(a-b)) is antisymmetric (fl(c*(-d)) = -fl(c*d) and truncation is
// symmetric about zero), so every flux is evaluated ONCE and applied with opposite signs to its two
// cells: a warp owns one block of species rows of one tile column and walks down the grid rows
// carrying, in registers, the current row's densities, its north fluxes and its horizontal net flux.
// Every density is therefore loaded from memory exactly once (16-byte vectors, 512 contiguous
// bytes per warp) and written once: 8 B of traffic per cell-species update.
#pragma once
#include "k_common.cuh"

#ifndef ECO_MIG_MINB
#define ECO_MIG_MINB 4
#endif
#ifndef ECO_MIG_PF
#define ECO_MIG_PF 0      // grid rows between an L2 prefetch and the register load of a row; 0 = none (measured: 4/8/16 rows ahead all SLOWER, 2.3-2.6 vs 1.85 ms on C3)
#endif

namespace eco {

struct MigParams {
  Layout L;                   // source buffer
  int32_t* dst;               // owned row 0 of the destination buffer (same layout)
  const int32_t* top_row;     // grid row y0-1 (ghost row or a peer GPU's boundary row), tile 0
  const int32_t* bottom_row;  // grid row y0+ny
  const double* cm;           // [cap] capped migration coefficient per species row (0 for padding rows)
  uint32_t nblocks;           // ceil(nsp / ROWS)
  uint32_t yseg, nseg;        // grid rows per work unit, units along y
  unsigned long long* total;
};

// 2^52 + n as a double: pairing the int with a constant high word costs no arithmetic
__device__ __forceinline__ double big(int n) { return __hiloint2double(0x43300000, n); }
// trunc(c * (a - b)) for non-negative int32 a, b: the difference of two doubles in [2^52, 2^53) is exact
__device__ __forceinline__ int flux(double c, double ba, double bb) {
  return __double2int_rz(__dmul_rn(c, __dadd_rn(ba, -bb)));
}

// Cells across the tile boundary.  A lane on the first (last) quad of its species row loads ONE extra cell per grid
// row: the last cell of the west tile (first cell of the east tile).  Where the grid has no such neighbour the
// extra load aliases the lane's own boundary cell, so the difference -- and the flux -- is 0 without any flag
// (open boundaries, reference :593-600).  Lanes in the middle of a row alias their own first cell (value unused).
struct MigRow {
  int4 v;      // this lane's 4 cells
  int e;       // west neighbour (first quad) / east neighbour (last quad)
  int e2;      // TW == 4 only (one lane is both first and last quad): east neighbour, `e` is the west one
};

template <int TW>
__device__ __forceinline__ MigRow mig_load(const int32_t* p, long long eoff, long long eoff2) {
  MigRow m;
  m.v = *reinterpret_cast<const int4*>(p);
  m.e = *(p + eoff);
  m.e2 = 0;
  if (TW == 4) m.e2 = *(p + eoff2);
  return m;
}

// horizontal net flux of the 4 cells held as 2^52+n doubles: hs_k = T(west_k - own_k) - T(own_k - east_k).
// RAG (last tile of a row wider than the grid): `rag` = 1 + number of real cells among the lane's four and its
// east neighbour; padding cells hold 0 and exchange nothing.
template <int TW, bool RAG>
__device__ __forceinline__ int4 mig_horizontal(const double (&b)[4], int e, int e2, double c, bool first_q, bool last_q,
                                               uint32_t rag) {
  int lw = __shfl_up_sync(FULL, __double2loint(b[3]), 1);
  if (first_q) lw = e;
  const int re = TW == 4 ? e2 : e;
  int t0 = flux(c, big(lw), b[0]), t1 = flux(c, b[0], b[1]), t2 = flux(c, b[1], b[2]), t3 = flux(c, b[2], b[3]);
  int te = flux(c, b[3], big(re));                       // west flux of the east tile's first cell
  if (RAG) {
    if (rag < 2) t0 = 0;
    if (rag < 3) t1 = 0;
    if (rag < 4) t2 = 0;
    if (rag < 5) t3 = 0;
    if (rag < 6) te = 0;
  }
  int tn = __shfl_down_sync(FULL, t0, 1);
  if (last_q) tn = te;
  return make_int4(t0 - t1, t1 - t2, t2 - t3, t3 - tn);
}

// State a warp carries down the grid rows: the current row as 2^52+n doubles (the low word IS the density),
// its extra boundary cell(s) and its north fluxes.
struct MigCarry {
  double b[4];
  int e, e2;
  int4 tn;
};

// Finish the current row given the row below it (`nxt`), store it, and make `nxt` the current row.
template <int TW, bool RAG>
__device__ __forceinline__ void mig_step(MigCarry& k, const MigRow& nxt, double c, bool first_q, bool last_q, uint32_t rag,
                                         int32_t* __restrict__ dp, unsigned long long& acc) {
  const double bn[4] = {big(nxt.v.x), big(nxt.v.y), big(nxt.v.z), big(nxt.v.w)};
  // north flux of the row below = minus the south flux of this row
  const int4 ts = make_int4(flux(c, k.b[0], bn[0]), flux(c, k.b[1], bn[1]), flux(c, k.b[2], bn[2]), flux(c, k.b[3], bn[3]));
  const int4 hs = mig_horizontal<TW, RAG>(k.b, k.e, k.e2, c, first_q, last_q, rag);
  const int d0 = k.tn.x - ts.x + hs.x, d1 = k.tn.y - ts.y + hs.y, d2 = k.tn.z - ts.z + hs.z, d3 = k.tn.w - ts.w + hs.w;
  *reinterpret_cast<int4*>(dp) = make_int4(__double2loint(k.b[0]) + d0, __double2loint(k.b[1]) + d1,
                                           __double2loint(k.b[2]) + d2, __double2loint(k.b[3]) + d3);
  acc += (unsigned long long)((uint32_t)abs(d0) + (uint32_t)abs(d1)) + (unsigned long long)((uint32_t)abs(d2) + (uint32_t)abs(d3));
  k.b[0] = bn[0]; k.b[1] = bn[1]; k.b[2] = bn[2]; k.b[3] = bn[3];
  k.e = nxt.e; k.e2 = nxt.e2; k.tn = ts;
}

// A work unit = one block of species rows x one tile column x n grid rows.  The rows of the unit are streamed as
// the sequence  north, row(0), ..., row(n-1), south  where `north` / `south` are the rows just outside the unit
// (ghost rows or a peer GPU's boundary rows at the slab ends) or, at the edge of the GLOBAL grid, the unit's own
// first / last row again (difference 0 -> flux 0).  Loads run two rows ahead of the arithmetic; the main loop
// (two rows per trip, no register rotation) has no boundary cases.
template <int TW, bool RAG>
__device__ __forceinline__ void mig_unit(const int32_t* __restrict__ p0, const int32_t* __restrict__ pn,
                                         const int32_t* __restrict__ ps, int32_t* __restrict__ dp, long long stride, uint32_t n,
                                         long long eoff, long long eoff2, double c, bool first_q, bool last_q, uint32_t rag,
                                         unsigned long long& acc) {
  auto rowp = [&](uint32_t r) -> const int32_t* { return r < n ? p0 + (long long)r * stride : ps; };
  const int4 up = *reinterpret_cast<const int4*>(pn);
  const MigRow first = mig_load<TW>(p0, eoff, eoff2);
  MigRow ra = mig_load<TW>(rowp(1), eoff, eoff2);
  MigRow rb = mig_load<TW>(rowp(2), eoff, eoff2);
  MigCarry k;
  k.b[0] = big(first.v.x); k.b[1] = big(first.v.y); k.b[2] = big(first.v.z); k.b[3] = big(first.v.w);
  k.e = first.e; k.e2 = first.e2;
  k.tn = make_int4(flux(c, big(up.x), k.b[0]), flux(c, big(up.y), k.b[1]), flux(c, big(up.z), k.b[2]),
                   flux(c, big(up.w), k.b[3]));                                             // T(n[y-1] - n[y])
  uint32_t i = 0;
  const int32_t* pl = p0 + 3 * stride;                                                      // row(i+3)
  for (; i + 4 < n; i += 2) {
    if (ECO_MIG_PF > 0 && i + 4 + ECO_MIG_PF < n) {   // warp-uniform; the registers only hold two rows in flight, L2 holds the rest
      asm volatile("prefetch.global.L2 [%0];" ::"l"(pl + (long long)ECO_MIG_PF * stride));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(pl + (long long)(ECO_MIG_PF + 1) * stride));
    }
    mig_step<TW, RAG>(k, ra, c, first_q, last_q, rag, dp, acc);
    ra = mig_load<TW>(pl, eoff, eoff2);
    mig_step<TW, RAG>(k, rb, c, first_q, last_q, rag, dp + stride, acc);
    rb = mig_load<TW>(pl + stride, eoff, eoff2);
    pl += 2 * stride;
    dp += 2 * stride;
  }
#pragma unroll 1
  for (; i < n; ++i) {                                                                      // at most 4 rows
    mig_step<TW, RAG>(k, ra, c, first_q, last_q, rag, dp, acc);
    ra = rb;
    rb = mig_load<TW>(rowp(i + 3), eoff, eoff2);
    dp += stride;
  }
}

template <int TW>
__global__ void __launch_bounds__(256, ECO_MIG_MINB) migrate_kernel(const MigParams p) {
  constexpr int R = LaneMap<TW>::ROWS, QPR = LaneMap<TW>::QPR;
  const Layout& L = p.L;
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  const uint32_t rr = LaneMap<TW>::row(lane), q = LaneMap<TW>::quad(lane);
  const unsigned long long nunits = (unsigned long long)p.nseg * L.ntx * p.nblocks;
  const long long stride = (long long)L.row_stride;
  const bool first_q = q == 0, last_q = q == QPR - 1;
  unsigned long long acc = 0;
  for (unsigned long long w = (unsigned long long)blockIdx.x * wpb + (threadIdx.x >> 5); w < nunits;
       w += (unsigned long long)gridDim.x * wpb) {
    const uint32_t b = (uint32_t)(w % p.nblocks);
    const uint32_t tx = (uint32_t)((w / p.nblocks) % L.ntx);
    const uint32_t seg = (uint32_t)(w / ((unsigned long long)p.nblocks * L.ntx));
    const uint32_t ys = seg * p.yseg, ye = min(L.ny, ys + p.yseg), n = ye - ys;
    const uint32_t row = b * R + rr;
    const double c = __ldg(p.cm + row);
    const uint32_t x0 = tx * TW + 4 * q;
    const bool has_l = x0 > 0, has_r = x0 + 4 < L.nx;
    // element offsets of the extra cell(s) relative to the lane's vector
    const long long eo_w = has_l ? (long long)(TW - 1) - (long long)L.tile_stride : 0;   // west tile's last cell | own cell 0
    const long long eo_e = has_r ? (long long)L.tile_stride - (long long)(TW - 4) : 3;   // east tile's first cell | own cell 3
    const long long eoff = TW == 4 ? eo_w : (first_q ? eo_w : (last_q ? eo_e : 0));
    const size_t lane_off = (size_t)tx * L.tile_stride + (size_t)row * TW + 4 * q;
    const int32_t* p0 = L.base + (size_t)ys * L.row_stride + lane_off;                      // row(0)
    const int32_t* pn = L.y0 + ys == 0 ? p0 : (ys == 0 ? p.top_row + lane_off : p0 - stride);
    const int32_t* ps = L.y0 + ye >= L.nyg ? p0 + (long long)(n - 1) * stride
                                           : (ye == L.ny ? p.bottom_row + lane_off : p0 + (long long)n * stride);
    int32_t* dp = p.dst + (size_t)ys * L.row_stride + lane_off;
    if ((tx + 1) * TW > L.nx) {   // warp-uniform
      const uint32_t rag = 1u + (x0 >= L.nx ? 0u : min(5u, L.nx - x0));
      mig_unit<TW, true>(p0, pn, ps, dp, stride, n, eoff, eo_e, c, first_q, last_q, rag, acc);
    } else {
      mig_unit<TW, false>(p0, pn, ps, dp, stride, n, eoff, eo_e, c, first_q, last_q, 0u, acc);
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
  if (lane == 0 && acc) atomicAdd(p.total, acc);
}

}  // namespace eco
