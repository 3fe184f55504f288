// k_migrate.cuh -- nearest-neighbour migration (reference SpatialModel::migrate,
// models/ecolab_model.cc:616-703):
//     delta(c,s) = sum over existing W,E,N,S neighbours b of (int)(cm[s] * (double)(n_b - n_c))   (:631-642)
//     n_new = n_old + delta  (synchronous: reads `src`, writes `dst`)                              (:649-664)
//     returns sum |delta|                                                                          (:660, :702)
// The pairwise term T(a,b) = trunc(cm*(a-b)) is antisymmetric (fl(c*(-d)) = -fl(c*d) and truncation is
// symmetric about zero), so every flux is evaluated ONCE and applied with opposite signs to its two
// cells: a warp owns one block of species rows of one tile column and walks down the grid rows
// carrying, in registers, the current row's densities and its north fluxes.  Every density is loaded
// from HBM exactly once and written once: 8 B of traffic per cell-species update.
//
// One kernel: migrate_kernel, the register walker -- each warp walks down the grid rows of one block of species rows of one
// tile column with five rows in flight in registers.  A tile-column pipeline through a shared-memory ring (bulk copies by a
// producer lane, 16 consumer warps; round 2) was bit-exact but slower at every ring depth (C3: 2.5 vs 1.57 ms,
// profiles/r2_variants.txt) and was removed.
#pragma once
#include "k_common.cuh"

namespace eco {

struct MigParams {
  Layout L;                   // source buffer
  int32_t* dst;               // owned row 0 of the destination buffer (same layout)
  const int32_t* top_row;     // grid row y0-1 (ghost row or a peer GPU's boundary row), tile 0
  const int32_t* bottom_row;  // grid row y0+ny
  const double* cm;           // [cap] capped migration coefficient per species row (0 for padding rows)
  uint32_t yseg, nseg;        // grid rows per work unit, units along y
  // multi-GPU, peer rows read over NVLink (null = ghost rows already filled): the neighbour ranks' "generate done"
  // epoch counters; a boundary unit's warp waits until the counter reaches `epoch` before it copies the peer's row
  const unsigned int* top_flag;
  const unsigned int* bottom_flag;
  unsigned int epoch;
  unsigned int* err;          // set to 1 if a flag wait timed out
  unsigned long long* total;
  unsigned int* next_unit;    // walker: work-unit counter, zero at launch
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
  int e;       // west neighbour (first lane of a species row) / east neighbour (last lane)
  int e2;      // DUALE only (one lane holds a whole species row): east neighbour, `e` is the west one
};

template <bool DUALE>
__device__ __forceinline__ MigRow mig_load(const int32_t* p, long long eoff, long long eoff2) {
  MigRow m;
  ECO_CHK(p, 16, 20u);
  ECO_CHK(p + eoff, 4, 21u);
  m.v = *reinterpret_cast<const int4*>(p);
  m.e = *(p + eoff);
  m.e2 = 0;
  if (DUALE) { ECO_CHK(p + eoff2, 4, 22u); m.e2 = *(p + eoff2); }
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

// State a warp carries down the grid rows: the current row as 2^52+n doubles (the low word IS the density) and its
// north fluxes.
struct MigState {
  double b[4];
  int4 tn;
};

__device__ __forceinline__ void mig_start(MigState& k, const int4 up, const int4 cur, double c) {
  k.b[0] = big(cur.x); k.b[1] = big(cur.y); k.b[2] = big(cur.z); k.b[3] = big(cur.w);
  k.tn = make_int4(flux(c, big(up.x), k.b[0]), flux(c, big(up.y), k.b[1]), flux(c, big(up.z), k.b[2]),
                   flux(c, big(up.w), k.b[3]));                                             // T(n[y-1] - n[y])
}

// Finish the current row given the row below it (`nv`) and the current row's extra cell(s), store it, and make `nv`
// the current row.
template <bool DUALE, bool RAG>
__device__ __forceinline__ void mig_finish(MigState& k, const int4 nv, int e, int e2, double c, bool first_q, bool last_q,
                                           uint32_t rag, int32_t* __restrict__ dp, unsigned long long& acc) {
  const double nb[4] = {big(nv.x), big(nv.y), big(nv.z), big(nv.w)};
  // north flux of the row below = minus the south flux of this row
  const int4 ts = make_int4(flux(c, k.b[0], nb[0]), flux(c, k.b[1], nb[1]), flux(c, k.b[2], nb[2]), flux(c, k.b[3], nb[3]));
  const int4 hs = mig_horizontal<DUALE, RAG>(k.b, e, e2, c, first_q, last_q, rag);
  const int d0 = k.tn.x - ts.x + hs.x, d1 = k.tn.y - ts.y + hs.y, d2 = k.tn.z - ts.z + hs.z, d3 = k.tn.w - ts.w + hs.w;
  const int4 out = make_int4(__double2loint(k.b[0]) + d0, __double2loint(k.b[1]) + d1, __double2loint(k.b[2]) + d2,
                             __double2loint(k.b[3]) + d3);
  ECO_CHK(dp, 16, 23u);
  __stcs(reinterpret_cast<int4*>(dp), out);   // streaming (evict-first): the row is not read again by this kernel
  acc += (unsigned long long)((uint32_t)abs(d0) + (uint32_t)abs(d1)) + (unsigned long long)((uint32_t)abs(d2) + (uint32_t)abs(d3));
  k.b[0] = nb[0]; k.b[1] = nb[1]; k.b[2] = nb[2]; k.b[3] = nb[3];
  k.tn = ts;
}

// =====================================================================================================================
// the register walker.  A work unit = one block of species rows x one tile x n grid rows, streamed as the
// sequence  north, row(0), ..., row(n-1), south  where `north` / `south` are the rows just outside the unit (ghost
// rows or a peer GPU's boundary rows at the slab ends) or, at the edge of the GLOBAL grid, the unit's own first / last
// row again (difference 0 -> flux 0).  NB rows are in flight per warp in registers; a row's load is re-issued the
// moment the row is unpacked.
// =====================================================================================================================
#ifndef ECO_MIG_THREADS
#define ECO_MIG_THREADS 256
#endif
#ifndef ECO_MIG_NB
#define ECO_MIG_NB 6   // rows in flight per warp; C4 / C3 ms with the dynamic unit order: 4: 24.6 / -, 5: 23.5 / 1.567, 6: 23.0 / 1.551, 7: 25.4 / 1.531 (spills), 8: 26.2 / 1.715
#endif
constexpr int kMigThreads = ECO_MIG_THREADS;
constexpr int kMigRowsInFlight = ECO_MIG_NB;   // (round 1, static unit order, C3: 2: 1.67  3: 1.73  4: 1.63  5: 1.60  6: 1.89  8: 2.04)

template <bool DUALE, bool RAG, int NB>
__device__ __forceinline__ void mig_unit(const int32_t* __restrict__ p0, const int32_t* __restrict__ pn,
                                         const int32_t* __restrict__ ps, int32_t* __restrict__ dp, long long stride, uint32_t n,
                                         long long eoff, long long eoff2, double c, bool first_q, bool last_q, uint32_t rag,
                                         unsigned long long& acc) {
  auto rowp = [&](uint32_t r) -> const int32_t* { return r < n ? p0 + (long long)r * stride : ps; };
  ECO_CHK(pn, 16, 24u);
  const int4 up = *reinterpret_cast<const int4*>(pn);
  const MigRow first = mig_load<DUALE>(p0, eoff, eoff2);
  MigRow r[NB];                                                                             // rows i+1 .. i+NB, in flight
#pragma unroll
  for (int j = 0; j < NB; ++j) r[j] = mig_load<DUALE>(rowp(1 + j), eoff, eoff2);
  MigState k;
  mig_start(k, up, first.v, c);
  int e = first.e, e2 = first.e2;
  uint32_t i = 0;
  const int32_t* pl = p0 + (long long)(NB + 1) * stride;                                    // row(i+NB+1)
  for (; i + 2 * NB < n; i += NB) {                                                         // NB rows per trip, no register rotation
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const MigRow nx = r[j];
      r[j] = mig_load<DUALE>(pl + (long long)j * stride, eoff, eoff2);
      mig_finish<DUALE, RAG>(k, nx.v, e, e2, c, first_q, last_q, rag, dp + (long long)j * stride, acc);
      e = nx.e; e2 = nx.e2;
    }
    pl += (long long)NB * stride;
    dp += (long long)NB * stride;
  }
#pragma unroll 1
  for (; i < n; ++i) {                                                                      // at most 2*NB rows
    const MigRow nx = r[0];
#pragma unroll
    for (int j = 0; j + 1 < NB; ++j) r[j] = r[j + 1];
    r[NB - 1] = mig_load<DUALE>(rowp(i + NB + 1), eoff, eoff2);
    mig_finish<DUALE, RAG>(k, nx.v, e, e2, c, first_q, last_q, rag, dp, acc);
    e = nx.e; e2 = nx.e2;
    dp += stride;
  }
}

// element offsets of a lane's extra cell(s) relative to its vector: west tile's last cell | own cell 0, east tile's
// first cell | own cell 3
__device__ __forceinline__ void mig_edge_offsets(const Layout& L, uint32_t x0, uint32_t tw, long long& eo_w, long long& eo_e) {
  const bool has_l = x0 > 0, has_r = x0 + 4 < L.nx;
  eo_w = has_l ? (long long)(tw - 1) - (long long)L.tile_stride : 0;
  eo_e = has_r ? (long long)L.tile_stride - (long long)(tw - 4) : 3;
}

// publish this rank's epoch counter to the neighbour ranks (they map it through CUDA IPC)
__global__ void flag_publish_kernel(unsigned int* flag, unsigned int value) {
  __threadfence_system();   // everything this stream did before is visible to the peers before the counter moves
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}

// poll a neighbour rank's "generate done" counter (peer memory over NVLink); bounded so that a lost peer cannot hang the GPU
__device__ __forceinline__ void mig_wait_flag(const unsigned int* flag, unsigned int epoch, unsigned int* err) {
  const long long t0 = clock64();
  for (;;) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if ((int)(v - epoch) >= 0) return;
    if (clock64() - t0 > 20000000000ll) { atomicExch(err, 1u); return; }   // ~10 s
    __nanosleep(200);
  }
}

template <int TW>
__global__ void __launch_bounds__(kMigThreads, 512 / kMigThreads) migrate_kernel(const MigParams p) {
  constexpr int QPR = TW / 4, RW = 32 / QPR;   // lanes per species row, species rows per warp
  constexpr bool DUALE = QPR == 1;
  const Layout& L = p.L;
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  const uint32_t rr = lane / QPR, q = lane % QPR;
  const uint32_t nrowblk = (L.nsp + RW - 1) / RW;
  const long long stride = (long long)L.row_stride;
  const bool first_q = q == 0, last_q = q == QPR - 1;
  unsigned long long acc = 0;
  // Work units are handed out dynamically, tile column fastest: unit u = (segment, group of `wpb` consecutive species-row
  // blocks, tile column) goes to the next free CTA, whose warps take one block each (a CTA reads wpb x 512 contiguous bytes per
  // grid row).  The CTAs running at any time therefore work on ADJACENT tile columns of the same rows, started within a
  // fraction of a microsecond of each other: the 32-byte sectors a warp reads from its neighbour tiles (one cell per
  // species row and side) are lines the neighbour CTA is streaming through L2 at that moment.  With the static
  // block-fastest order of round 1 a wave covered ~18 tile columns and drifted apart: at 4096 x 4096 x 1024 the kernel
  // read 80-82 GB from DRAM for 68.7 GB of densities (ncu, profiles/r2_c4_traffic.csv).
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t ngr = (nrowblk + wpb - 1) / wpb;
  const unsigned long long ncu = (unsigned long long)p.nseg * L.ntx * ngr;
  __shared__ unsigned int s_unit;
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) s_unit = atomicAdd(p.next_unit, 1u);
    __syncthreads();
    const unsigned long long w = s_unit;
    if (w >= ncu) break;
    const uint32_t tx = (uint32_t)(w % L.ntx);
    const uint32_t b = (uint32_t)((w / L.ntx) % ngr) * wpb + warp;
    const uint32_t seg = (uint32_t)(w / ((unsigned long long)ngr * L.ntx));
    if (b >= nrowblk) continue;
    const uint32_t ys = seg * p.yseg, ye = min(L.ny, ys + p.yseg), n = ye - ys;
    const uint32_t row = min(b * RW + rr, L.cap - 1);      // cap is a multiple of 32: the clamp never triggers, it bounds the address
    const double c = __ldg(p.cm + row);
    const uint32_t x0 = tx * TW + 4 * q;
    long long eo_w, eo_e;
    mig_edge_offsets(L, x0, TW, eo_w, eo_e);
    const long long eoff = (DUALE || first_q) ? eo_w : (last_q ? eo_e : 0);
    const size_t lane_off = (size_t)tx * L.tile_stride + (size_t)row * TW + 4 * q;
    const int32_t* p0 = L.base + (size_t)ys * L.row_stride + lane_off;                      // row(0)
    const int32_t* pn = L.y0 + ys == 0 ? p0 : (ys == 0 ? p.top_row + lane_off : p0 - stride);
    const int32_t* ps = L.y0 + ye >= L.nyg ? p0 + (long long)(n - 1) * stride
                                           : (ye == L.ny ? p.bottom_row + lane_off : p0 + (long long)n * stride);
    int32_t* dp = p.dst + (size_t)ys * L.row_stride + lane_off;
    // peer-memory halo: a unit that touches a neighbour rank's row first waits for that rank's epoch counter
    if ((p.top_flag && ys == 0 && L.y0 != 0) || (p.bottom_flag && ye == L.ny && L.y0 + ye < L.nyg)) {
      if (lane == 0) {
        if (p.top_flag && ys == 0 && L.y0 != 0) mig_wait_flag(p.top_flag, p.epoch, p.err);
        if (p.bottom_flag && ye == L.ny && L.y0 + ye < L.nyg) mig_wait_flag(p.bottom_flag, p.epoch, p.err);
      }
      __syncwarp();
    }
    const bool ragged = (tx + 1) * TW > L.nx;      // warp-uniform: the ragged last tile
    const uint32_t rag = 1u + (x0 >= L.nx ? 0u : min(5u, L.nx - x0));
    if (ragged) mig_unit<DUALE, true, kMigRowsInFlight>(p0, pn, ps, dp, stride, n, eoff, eo_e, c, first_q, last_q, rag, acc);
    else mig_unit<DUALE, false, kMigRowsInFlight>(p0, pn, ps, dp, stride, n, eoff, eo_e, c, first_q, last_q, 0u, acc);
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
  if (lane == 0 && acc) atomicAdd(p.total, acc);
}

}  // namespace eco
