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
#define ECO_MIG_MINB 2     // resident CTAs per SM the register budget is set for (2 x 256 threads -> up to 128 registers)
#endif
#ifndef ECO_MIG_NB
#define ECO_MIG_NB 5       // grid rows a warp keeps in flight in registers.  Measured on C3 (ms): NB/CTAs per SM =
                           // 2/4: 2.03 (spills)  2/3: 1.68  3/3: 2.03 (spills)  3/2: 1.73  4/2: 1.63  5/2: 1.61  6/2: 1.89  8/1: 2.04
#endif
#ifndef ECO_MIG_ALIAS
#define ECO_MIG_ALIAS 1    // extra-cell loads unpredicated (measured on C3: 1.85 ms, predicated 1.96 ms)
#endif
#ifndef ECO_MIG_STCS
#define ECO_MIG_STCS 1     // 1: streaming (evict-first) stores (with 5 rows in flight: 1.60 vs 1.61 ms)
#endif
#ifndef ECO_MIG_LDCS
#define ECO_MIG_LDCS 0     // 1: streaming loads of the lane's vector
#endif
#ifndef ECO_MIG_THREADS
#define ECO_MIG_THREADS 256
#endif
#ifndef ECO_MIG_PF
#define ECO_MIG_PF 0      // grid rows between an L2 prefetch and the register load of a row; 0 = none (per-lane prefetches in the earlier kernel: 4/8/16 rows ahead all SLOWER, 2.3-2.6 vs 1.85 ms on C3); now counted in trips of ECO_MIG_NB rows, bulk prefetch per warp row
#endif

namespace eco {

struct MigParams {
  Layout L;                   // source buffer
  int32_t* dst;               // owned row 0 of the destination buffer (same layout)
  const int32_t* top_row;     // grid row y0-1 (ghost row or a peer GPU's boundary row), tile 0
  const int32_t* bottom_row;  // grid row y0+ny
  const double* cm;           // [cap] capped migration coefficient per species row (0 for padding rows)
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
// (open boundaries, reference :593-600).  Lanes in the middle of a row load nothing extra.
struct MigRow {
  int4 v;      // this lane's 4 cells
  int e;       // west neighbour (first lane of a species row) / east neighbour (last lane)
  int e2;      // DUALE only (one lane holds a whole species row of the warp's span): east neighbour, `e` is the west one
};

template <bool DUALE>
__device__ __forceinline__ MigRow mig_load(const int32_t* p, long long eoff, long long eoff2, bool has_e) {
  MigRow m;
#if ECO_MIG_LDCS
  m.v = __ldcs(reinterpret_cast<const int4*>(p));
#else
  m.v = *reinterpret_cast<const int4*>(p);
#endif
  m.e = 0; m.e2 = 0;
#if ECO_MIG_ALIAS
  m.e = *(p + eoff);                 // every lane loads: lanes in the middle of a row alias their own first cell
#else
  if (has_e) m.e = *(p + eoff);
#endif
  if (DUALE) m.e2 = *(p + eoff2);
  return m;
}

// horizontal net flux of the 4 cells held as 2^52+n doubles: hs_k = T(west_k - own_k) - T(own_k - east_k).
// RAG (last tile of a row wider than the grid): `rag` = 1 + number of real cells among the lane's four and its
// east neighbour; padding cells hold 0 and exchange nothing.
template <bool DUALE, bool RAG>
__device__ __forceinline__ int4 mig_horizontal(const double (&b)[4], int e, int e2, double c, bool first_q, bool last_q,
                                               uint32_t rag) {
  int lw = __shfl_up_sync(FULL, __double2loint(b[3]), 1);
  if (first_q) lw = e;
  const int re = DUALE ? e2 : e;
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

// A freshly loaded row unpacked into the carry's form.  Unpacking is the LAST use of the load's registers, so the
// caller re-issues the load for a later row right after it and before the row's arithmetic: every load then has
// two full rows of arithmetic (its own buffer's and the other buffer's) to arrive, not one.
struct MigNext {
  double b[4];
  int e, e2;
};
__device__ __forceinline__ MigNext mig_unpack(const MigRow& r) {
  MigNext x;
  x.b[0] = big(r.v.x); x.b[1] = big(r.v.y); x.b[2] = big(r.v.z); x.b[3] = big(r.v.w);
  x.e = r.e; x.e2 = r.e2;
  return x;
}

// Finish the current row given the row below it (`nx`), store it, and make `nx` the current row.
template <bool DUALE, bool RAG>
__device__ __forceinline__ void mig_finish(MigCarry& k, const MigNext& nx, double c, bool first_q, bool last_q, uint32_t rag,
                                           int32_t* __restrict__ dp, unsigned long long& acc) {
#ifdef ECO_MIG_ABLATE   // cost attribution only (wrong results on purpose): the same loads and stores, no flux arithmetic
  {
    const int4 o = make_int4(__double2loint(k.b[0]) + nx.e, __double2loint(k.b[1]), __double2loint(k.b[2]), __double2loint(k.b[3]) + k.e);
    *reinterpret_cast<int4*>(dp) = o;
    k.b[0] = nx.b[0]; k.b[1] = nx.b[1]; k.b[2] = nx.b[2]; k.b[3] = nx.b[3];
    k.e = nx.e;
    return;
  }
#endif
  // north flux of the row below = minus the south flux of this row
  const int4 ts = make_int4(flux(c, k.b[0], nx.b[0]), flux(c, k.b[1], nx.b[1]), flux(c, k.b[2], nx.b[2]), flux(c, k.b[3], nx.b[3]));
  const int4 hs = mig_horizontal<DUALE, RAG>(k.b, k.e, k.e2, c, first_q, last_q, rag);
  const int d0 = k.tn.x - ts.x + hs.x, d1 = k.tn.y - ts.y + hs.y, d2 = k.tn.z - ts.z + hs.z, d3 = k.tn.w - ts.w + hs.w;
  const int4 out = make_int4(__double2loint(k.b[0]) + d0, __double2loint(k.b[1]) + d1, __double2loint(k.b[2]) + d2,
                             __double2loint(k.b[3]) + d3);
#if ECO_MIG_STCS
  __stcs(reinterpret_cast<int4*>(dp), out);
#else
  *reinterpret_cast<int4*>(dp) = out;
#endif
  acc += (unsigned long long)((uint32_t)abs(d0) + (uint32_t)abs(d1)) + (unsigned long long)((uint32_t)abs(d2) + (uint32_t)abs(d3));
  k.b[0] = nx.b[0]; k.b[1] = nx.b[1]; k.b[2] = nx.b[2]; k.b[3] = nx.b[3];
  k.e = nx.e; k.e2 = nx.e2; k.tn = ts;
}
template <bool DUALE, bool RAG>
__device__ __forceinline__ void mig_step(MigCarry& k, const MigRow& nxt, double c, bool first_q, bool last_q, uint32_t rag,
                                         int32_t* __restrict__ dp, unsigned long long& acc) {
  mig_finish<DUALE, RAG>(k, mig_unpack(nxt), c, first_q, last_q, rag, dp, acc);
}

// A work unit = one block of species rows x one tile column x n grid rows.  The rows of the unit are streamed as
// the sequence  north, row(0), ..., row(n-1), south  where `north` / `south` are the rows just outside the unit
// (ghost rows or a peer GPU's boundary rows at the slab ends) or, at the edge of the GLOBAL grid, the unit's own
// first / last row again (difference 0 -> flux 0).  Loads run two rows ahead of the arithmetic (see MigNext); the
// main loop (two rows per trip, no register rotation) has no boundary cases.
template <bool DUALE, bool RAG, int NB>
__device__ __forceinline__ void mig_unit(const int32_t* __restrict__ p0, const int32_t* __restrict__ pn,
                                         const int32_t* __restrict__ ps, int32_t* __restrict__ dp, long long stride, uint32_t n,
                                         long long eoff, long long eoff2, bool has_e, double c, bool first_q, bool last_q,
                                         uint32_t rag, unsigned long long& acc) {
  auto rowp = [&](uint32_t r) -> const int32_t* { return r < n ? p0 + (long long)r * stride : ps; };
  const int4 up = *reinterpret_cast<const int4*>(pn);
  const MigRow first = mig_load<DUALE>(p0, eoff, eoff2, has_e);
  MigRow r[NB];                                                                             // rows i+1 .. i+NB, in flight
#pragma unroll
  for (int j = 0; j < NB; ++j) r[j] = mig_load<DUALE>(rowp(1 + j), eoff, eoff2, has_e);
  MigCarry k;
  k.b[0] = big(first.v.x); k.b[1] = big(first.v.y); k.b[2] = big(first.v.z); k.b[3] = big(first.v.w);
  k.e = first.e; k.e2 = first.e2;
  k.tn = make_int4(flux(c, big(up.x), k.b[0]), flux(c, big(up.y), k.b[1]), flux(c, big(up.z), k.b[2]),
                   flux(c, big(up.w), k.b[3]));                                             // T(n[y-1] - n[y])
  uint32_t i = 0;
  const int32_t* pl = p0 + (long long)(NB + 1) * stride;                                    // row(i+NB+1)
#if ECO_MIG_PF > 0   // experiment (warp span T = 1 only): the warp's 512-byte rows requested into L2 ECO_MIG_PF trips ahead, one bulk prefetch per row
  const uint32_t pf_lane = threadIdx.x & 31u;
  const int32_t* wb = p0 - 4 * pf_lane;
  if (pf_lane < ECO_MIG_PF * NB && NB + 1 + pf_lane < n) bulk_prefetch_l2(wb + (long long)(NB + 1 + pf_lane) * stride, 512u);
#endif
  for (; i + 2 * NB < n; i += NB) {                                                         // NB rows per trip, no register rotation
#if ECO_MIG_PF > 0
    if (pf_lane < NB) {
      const uint32_t pr = i + NB + 1 + ECO_MIG_PF * NB + pf_lane;
      if (pr < n) bulk_prefetch_l2(wb + (long long)pr * stride, 512u);
    }
#endif
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const MigNext nx = mig_unpack(r[j]);
      r[j] = mig_load<DUALE>(pl + (long long)j * stride, eoff, eoff2, has_e);
      mig_finish<DUALE, RAG>(k, nx, c, first_q, last_q, rag, dp + (long long)j * stride, acc);
    }
    pl += (long long)NB * stride;
    dp += (long long)NB * stride;
  }
#pragma unroll 1
  for (; i < n; ++i) {                                                                      // at most 2*NB rows
    const MigNext nx = mig_unpack(r[0]);
#pragma unroll
    for (int j = 0; j + 1 < NB; ++j) r[j] = r[j + 1];
    r[NB - 1] = mig_load<DUALE>(rowp(i + NB + 1), eoff, eoff2, has_e);
    mig_finish<DUALE, RAG>(k, nx, c, first_q, last_q, rag, dp, acc);
    dp += stride;
  }
}

// ---- the same unit with a per-warp shared-memory ring instead of the two register rows -------------------------------
// Each lane copies its 16 bytes (+ its extra cell) of the next ECO_MIG_RING grid rows into a private slot with cp.async
// (LDGSTS: no registers held while the load is in flight) and reads it back when the row's turn comes.  Bytes in flight
// per warp: ECO_MIG_RING x 640 instead of 2 x 640.  A lane only ever touches its own slot bytes: no barrier, only
// cp.async.wait_group.  Slot = 512 B of vectors + 128 B of extra cells.
#ifndef ECO_MIG_RING
#define ECO_MIG_RING 0
#endif
constexpr uint32_t kMigSlotBytes = 640;

__device__ __forceinline__ void mig_cp_row(uint32_t sv, uint32_t se, const int32_t* p, long long eoff, bool has_e) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sv), "l"(p) : "memory");
  if (has_e) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(se), "l"(p + eoff) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
}

template <bool RAG, int D>
__device__ __forceinline__ void mig_unit_ring(const int32_t* __restrict__ p0, const int32_t* __restrict__ pn,
                                              const int32_t* __restrict__ ps, int32_t* __restrict__ dp, long long stride,
                                              uint32_t n, long long eoff, bool has_e, double c, bool first_q, bool last_q,
                                              uint32_t rag, uint32_t sv0, uint32_t se0, unsigned long long& acc) {
  auto rowp = [&](uint32_t r) -> const int32_t* { return r < n ? p0 + (long long)r * stride : ps; };
#pragma unroll
  for (int k = 0; k < D; ++k) mig_cp_row(sv0 + k * kMigSlotBytes, se0 + k * kMigSlotBytes, rowp(k + 1), eoff, has_e);
  const int4 up = *reinterpret_cast<const int4*>(pn);
  const MigRow first = mig_load<false>(p0, eoff, 0, has_e);
  MigCarry k;
  k.b[0] = big(first.v.x); k.b[1] = big(first.v.y); k.b[2] = big(first.v.z); k.b[3] = big(first.v.w);
  k.e = first.e; k.e2 = 0;
  k.tn = make_int4(flux(c, big(up.x), k.b[0]), flux(c, big(up.y), k.b[1]), flux(c, big(up.z), k.b[2]),
                   flux(c, big(up.w), k.b[3]));                                             // T(n[y-1] - n[y])
  const int32_t* pl = rowp(D + 1);
  uint32_t so = 0;                                                                          // byte offset of the slot of row(i+1)
  for (uint32_t i = 0; i < n; ++i) {
    asm volatile("cp.async.wait_group %0;" ::"n"(D - 1) : "memory");
    MigRow nxt;
    asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(nxt.v.x), "=r"(nxt.v.y), "=r"(nxt.v.z), "=r"(nxt.v.w) : "r"(sv0 + so) : "memory");
    nxt.e = 0; nxt.e2 = 0;
    if (has_e) asm volatile("ld.shared.s32 %0, [%1];" : "=r"(nxt.e) : "r"(se0 + so) : "memory");
    mig_step<false, RAG>(k, nxt, c, first_q, last_q, rag, dp, acc);
    mig_cp_row(sv0 + so, se0 + so, pl, eoff, has_e);                                        // row(i+1+D), clamped to `south`
    pl = i + 2 + D < n ? pl + stride : ps;
    so = so + kMigSlotBytes == D * kMigSlotBytes ? 0u : so + kMigSlotBytes;
    dp += stride;
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// Lane map: a warp spans T adjacent tiles of one grid row: lane = (species row * T + tile) * QPR + quad, i.e. the
// LPR = T*QPR lanes of one species row hold 4*LPR CONSECUTIVE cells, neighbours across a tile boundary inside the span
// are one shuffle away, and only the two end lanes of a row load an extra cell (one 32-byte sector each).  T = 1 is
// the generate kernel's map (512 contiguous bytes per warp); larger T trades contiguity (32/T rows x 64 B per tile) for
// fewer extra sectors: per warp-row 2*32/LPR of them.  The host picks T (p.L.ntx must be a multiple of it).
template <int TW, int T>
__global__ void __launch_bounds__(ECO_MIG_THREADS, ECO_MIG_MINB) migrate_kernel(const MigParams p) {
  constexpr int QPR = TW / 4, LPR = QPR * T, RW = 32 / LPR;   // lanes per species row, species rows per warp
  constexpr bool DUALE = LPR == 1;
  static_assert(LPR <= 32 && 32 % LPR == 0, "span wider than a warp");
  const Layout& L = p.L;
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  const uint32_t rr = lane / LPR, t = (lane % LPR) / QPR, q = lane % QPR;
  const uint32_t ngroups = L.ntx / T, nrowblk = (L.nsp + RW - 1) / RW;
  const unsigned long long nunits = (unsigned long long)p.nseg * ngroups * nrowblk;
  const long long stride = (long long)L.row_stride;
  const bool first_q = lane % LPR == 0, last_q = lane % LPR == LPR - 1;
  const bool has_e = first_q || last_q;
#if ECO_MIG_RING > 0
  extern __shared__ __align__(16) unsigned char mig_smem[];
  const uint32_t sv0 = smem_u32(mig_smem) + (threadIdx.x >> 5) * (ECO_MIG_RING * kMigSlotBytes) + lane * 16u;
  const uint32_t se0 = sv0 - lane * 16u + 512u + lane * 4u;
#endif
  unsigned long long acc = 0;
  for (unsigned long long w = (unsigned long long)blockIdx.x * wpb + (threadIdx.x >> 5); w < nunits;
       w += (unsigned long long)gridDim.x * wpb) {
    const uint32_t b = (uint32_t)(w % nrowblk);
    const uint32_t g = (uint32_t)((w / nrowblk) % ngroups);
    const uint32_t seg = (uint32_t)(w / ((unsigned long long)nrowblk * ngroups));
    const uint32_t ys = seg * p.yseg, ye = min(L.ny, ys + p.yseg), n = ye - ys;
    const uint32_t row = min(b * RW + rr, L.cap - 1);      // cap is a multiple of 32: the clamp never triggers, it bounds the address
    const uint32_t tx = g * T + t;
    const double c = __ldg(p.cm + row);
    const uint32_t x0 = tx * TW + 4 * q;
    const bool has_l = x0 > 0, has_r = x0 + 4 < L.nx;
    // element offsets of the extra cell(s) relative to the lane's vector
    const long long eo_w = has_l ? (long long)(TW - 1) - (long long)L.tile_stride : 0;   // west tile's last cell | own cell 0
    const long long eo_e = has_r ? (long long)L.tile_stride - (long long)(TW - 4) : 3;   // east tile's first cell | own cell 3
    const long long eoff = (DUALE || first_q) ? eo_w : (last_q ? eo_e : 0);
    const size_t lane_off = (size_t)tx * L.tile_stride + (size_t)row * TW + 4 * q;
    const int32_t* p0 = L.base + (size_t)ys * L.row_stride + lane_off;                      // row(0)
    const int32_t* pn = L.y0 + ys == 0 ? p0 : (ys == 0 ? p.top_row + lane_off : p0 - stride);
    const int32_t* ps = L.y0 + ye >= L.nyg ? p0 + (long long)(n - 1) * stride
                                           : (ye == L.ny ? p.bottom_row + lane_off : p0 + (long long)n * stride);
    int32_t* dp = p.dst + (size_t)ys * L.row_stride + lane_off;
    const bool ragged = (g + 1) * T * TW > L.nx;   // warp-uniform: the span contains the ragged last tile
    const uint32_t rag = 1u + (x0 >= L.nx ? 0u : min(5u, L.nx - x0));
#if ECO_MIG_RING > 0
    if (!DUALE) {
      if (ragged) mig_unit_ring<true, ECO_MIG_RING>(p0, pn, ps, dp, stride, n, eoff, has_e, c, first_q, last_q, rag, sv0, se0, acc);
      else mig_unit_ring<false, ECO_MIG_RING>(p0, pn, ps, dp, stride, n, eoff, has_e, c, first_q, last_q, 0u, sv0, se0, acc);
      continue;
    }
#endif
    if (ragged) mig_unit<DUALE, true, ECO_MIG_NB>(p0, pn, ps, dp, stride, n, eoff, eo_e, has_e, c, first_q, last_q, rag, acc);
    else mig_unit<DUALE, false, ECO_MIG_NB>(p0, pn, ps, dp, stride, n, eoff, eo_e, has_e, c, first_q, last_q, 0u, acc);
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
  if (lane == 0 && acc) atomicAdd(p.total, acc);
}

}  // namespace eco
