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

struct MigRow {
  int4 v;      // this lane's 4 cells
  int l, r;    // neighbour cells across the tile boundary (only meaningful on the first / last quad)
};

template <int TW>
__device__ __forceinline__ MigRow mig_load(const int32_t* rowbase, size_t lane_off, size_t tile_stride, bool edge_l,
                                           bool edge_r) {
  MigRow m;
  const int32_t* p = rowbase + lane_off;
  m.v = *reinterpret_cast<const int4*>(p);
  m.l = 0; m.r = 0;
  if (edge_l) m.l = *(p - tile_stride + (TW - 1));   // last cell of the west tile, same species row (p is at quad 0)
  if (edge_r) m.r = *(p + tile_stride - (TW - 4));   // first cell of the east tile (p is at the last quad)
  return m;
}

// horizontal net flux of the 4 cells: hs_k = T(west_k - own_k) - T(own_k - east_k)
template <int TW>
__device__ __forceinline__ int4 mig_horizontal(const MigRow& m, double c, uint32_t q, bool has_l, bool has_r,
                                               bool ragged, uint32_t x0, uint32_t nx) {
  constexpr int QPR = LaneMap<TW>::QPR;
  int4 v = m.v;
  int lw = __shfl_up_sync(FULL, v.w, 1);
  if (q == 0) lw = m.l;
  if (!has_l) lw = v.x;                                  // no west neighbour: zero difference
  const double b0 = big(v.x), b1 = big(v.y), b2 = big(v.z), b3 = big(v.w);
  int t0 = flux(c, big(lw), b0), t1 = flux(c, b0, b1), t2 = flux(c, b1, b2), t3 = flux(c, b2, b3);
  const int re = has_r ? m.r : v.w;
  int te = flux(c, b3, big(re));                         // west flux of the east tile's first cell
  if (ragged) {                                          // last tile of a row wider than the grid: padding cells hold 0
    if (x0 >= nx) t0 = 0;
    if (x0 + 1 >= nx) t1 = 0;
    if (x0 + 2 >= nx) t2 = 0;
    if (x0 + 3 >= nx) t3 = 0;
    if (x0 + 4 >= nx) te = 0;
  }
  int tn = __shfl_down_sync(FULL, t0, 1);
  if (q == QPR - 1) tn = te;
  return make_int4(t0 - t1, t1 - t2, t2 - t3, t3 - tn);
}

__device__ __forceinline__ int4 mig_vertical(const int4& up, const int4& own, double c) {   // T(n[y-1] - n[y])
  return make_int4(flux(c, big(up.x), big(own.x)), flux(c, big(up.y), big(own.y)), flux(c, big(up.z), big(own.z)),
                   flux(c, big(up.w), big(own.w)));
}

template <int TW>
__global__ void __launch_bounds__(256) migrate_kernel(const MigParams p) {
  constexpr int R = LaneMap<TW>::ROWS, QPR = LaneMap<TW>::QPR;
  const Layout& L = p.L;
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  const uint32_t rr = LaneMap<TW>::row(lane), q = LaneMap<TW>::quad(lane);
  const unsigned long long nunits = (unsigned long long)p.nseg * L.ntx * p.nblocks;
  unsigned long long acc = 0;
  for (unsigned long long w = (unsigned long long)blockIdx.x * wpb + (threadIdx.x >> 5); w < nunits;
       w += (unsigned long long)gridDim.x * wpb) {
    const uint32_t b = (uint32_t)(w % p.nblocks);
    const uint32_t tx = (uint32_t)((w / p.nblocks) % L.ntx);
    const uint32_t seg = (uint32_t)(w / ((unsigned long long)p.nblocks * L.ntx));
    const uint32_t ys = seg * p.yseg, ye = min(L.ny, ys + p.yseg);
    const uint32_t row = b * R + rr;
    const double c = __ldg(p.cm + row);
    const uint32_t x0 = tx * TW + 4 * q;
    const bool has_l = x0 > 0, has_r = x0 + 4 < L.nx;
    const bool edge_l = q == 0 && tx > 0, edge_r = q == QPR - 1 && tx + 1 < L.ntx;
    const bool ragged = (tx + 1) * TW > L.nx;
    const size_t lane_off = (size_t)tx * L.tile_stride + (size_t)row * TW + 4 * q;
    auto rowbase = [&](int yl) -> const int32_t* {
      if (yl < 0) return p.top_row;
      if (yl >= (int)L.ny) return p.bottom_row;
      return L.base + (size_t)yl * L.row_stride;
    };
    // prologue: north flux of the first row, the row itself and its horizontal net
    MigRow cur = mig_load<TW>(rowbase((int)ys), lane_off, L.tile_stride, edge_l, edge_r);
    int4 tn = make_int4(0, 0, 0, 0);
    if (L.y0 + ys > 0) {
      const int4 up = *reinterpret_cast<const int4*>(rowbase((int)ys - 1) + lane_off);
      tn = mig_vertical(up, cur.v, c);
    }
    int4 hs = mig_horizontal<TW>(cur, c, q, has_l, has_r, ragged, x0, L.nx);
    bool has_nxt = L.y0 + ys + 1 < L.nyg;
    MigRow nxt = cur;
    if (has_nxt) nxt = mig_load<TW>(rowbase((int)ys + 1), lane_off, L.tile_stride, edge_l, edge_r);
    int32_t* dp = p.dst + (size_t)ys * L.row_stride + lane_off;
    for (uint32_t y = ys; y < ye; ++y) {
      // issue the loads of row y+2 before the arithmetic on row y+1 (register double buffering)
      const bool has_nn = y + 2 <= ye && L.y0 + y + 2 < L.nyg && y + 1 < ye;
      MigRow nn = nxt;
      if (has_nn) nn = mig_load<TW>(rowbase((int)y + 2), lane_off, L.tile_stride, edge_l, edge_r);
      int4 ts = make_int4(0, 0, 0, 0), hn = hs;
      if (has_nxt) {
        ts = mig_vertical(cur.v, nxt.v, c);          // north flux of row y+1 = minus the south flux of row y
        if (y + 1 < ye) hn = mig_horizontal<TW>(nxt, c, q, has_l, has_r, ragged, x0, L.nx);
      }
      const int d0 = tn.x - ts.x + hs.x, d1 = tn.y - ts.y + hs.y, d2 = tn.z - ts.z + hs.z, d3 = tn.w - ts.w + hs.w;
      *reinterpret_cast<int4*>(dp) = make_int4(cur.v.x + d0, cur.v.y + d1, cur.v.z + d2, cur.v.w + d3);
      acc += (unsigned long long)(uint32_t)abs(d0) + (uint32_t)abs(d1) + (uint32_t)abs(d2) + (uint32_t)abs(d3);
      dp += L.row_stride;
      cur = nxt; tn = ts; hs = hn; nxt = nn;
      has_nxt = L.y0 + y + 2 < L.nyg;
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
  if (lane == 0 && acc) atomicAdd(p.total, acc);
}

}  // namespace eco
