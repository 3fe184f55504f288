// k_misc.cuh -- condense / mutate / observables / layout-conversion kernels.  These are the
// streaming int32 passes around the hot loop; they take the tile width at run time.
#pragma once
#include "k_common.cuh"

namespace eco {

__device__ __forceinline__ size_t cell_addr(const Layout& L, uint32_t yl, uint32_t x, uint32_t s) {
  return ((size_t)yl * L.ntx + x / L.tw) * L.tile_stride + (size_t)s * L.tw + x % L.tw;
}

// ---------------------------------------------------------------------------------------------
// condense (reference SpatialModel::condense :171-196, EcolabPoint::condense :124-137)
// ---------------------------------------------------------------------------------------------
// mask[s] = 1 iff some owned cell has n != 0 (benign race: all writers store 1, as in the reference :176).
// A CTA first records what it sees in a shared-memory copy of the mask and only stores a flag that is not yet set there:
// once every living species has been seen (after the first tile or two) the pass is pure 16-byte loads.
__global__ void __launch_bounds__(256) alive_kernel(const Layout L, uint8_t* mask) {
  extern __shared__ uint8_t s_alive[];                 // [nsp]
  for (uint32_t i = threadIdx.x; i < L.nsp; i += blockDim.x) s_alive[i] = 0;
  __syncthreads();
  const uint32_t ntiles = L.ntx * L.ny, qpr = L.tw / 4, per = L.nsp * qpr;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int4* t = reinterpret_cast<const int4*>(L.base + (size_t)tile * L.tile_stride);
#pragma unroll 4
    for (uint32_t i = threadIdx.x; i < per; i += blockDim.x) {
      const int4 v = __ldcs(t + i);
      if (v.x | v.y | v.z | v.w) {
        const uint32_t sp = i / qpr;
        if (!s_alive[sp]) s_alive[sp] = 1;
      }
    }
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < L.nsp; i += blockDim.x)
    if (s_alive[i]) mask[i] = 1;
}

// in-place, order-preserving row compaction inside each tile (ghost rows included); map[s] = new index of kept row s.
// One warp per tile, 512 contiguous bytes (128/tw species rows) per step: all rows of a step are in registers before any
// is stored, and map[s] <= s, so a store never lands on a row that has not been read yet.  No block barriers.
__global__ void __launch_bounds__(256) compact_kernel(int32_t* buf, size_t tile_stride, uint32_t tw, uint32_t ntiles,
                                                      uint32_t nsp_old, uint32_t nkeep, const uint8_t* __restrict__ keep,
                                                      const uint32_t* __restrict__ map) {
  const uint32_t lane = threadIdx.x & 31u, wpb = blockDim.x >> 5;
  const uint32_t qpr = tw / 4, rows = 32 / qpr;         // lanes per species row, species rows per warp step
  const uint32_t rr = lane / qpr, q = lane % qpr;
  constexpr int U = 4;                                  // steps in flight
  for (uint32_t tile = blockIdx.x * wpb + (threadIdx.x >> 5); tile < ntiles; tile += gridDim.x * wpb) {
    int32_t* t = buf + (size_t)tile * tile_stride + 4 * q;
    for (uint32_t s0 = 0; s0 < nsp_old; s0 += rows * U) {
      int4 v[U];
      uint32_t dst[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t sp = s0 + u * rows + rr;
        dst[u] = 0xffffffffu;
        if (sp < nsp_old) {
          ECO_CHK(t + (size_t)sp * tw, 16, 30u);
          v[u] = *reinterpret_cast<const int4*>(t + (size_t)sp * tw);
          if (__ldg(keep + sp)) dst[u] = __ldg(map + sp);
        }
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (dst[u] != 0xffffffffu) { ECO_CHK(t + (size_t)dst[u] * tw, 16, 31u); *reinterpret_cast<int4*>(t + (size_t)dst[u] * tw) = v[u]; }
      __syncwarp();
    }
    for (uint32_t sp = nkeep + rr; sp < nsp_old; sp += rows) { ECO_CHK(t + (size_t)sp * tw, 16, 32u); *reinterpret_cast<int4*>(t + (size_t)sp * tw) = make_int4(0, 0, 0, 0); }
  }
}

// ---------------------------------------------------------------------------------------------
// mutate, cell side (reference EcolabPoint::mutate, models/ecolab_model.cc:278-314):
//   k = ROUND(ms[s] * n[c][s]) with the stage-1 Philox stream; n -= k; record (cell, slot, k) for k > 0
// The reference concatenates the per-cell lists in cell order and, inside a cell, in slot order (gen_index,
// include/arrays.h:2477-2486; its device build does this with joint_exclusive_scan, :284-306).  Here: two passes over
// the densities with a scan between them, no global atomics, no sort --
//   mutate_count_kernel   per-cell event counts (the draws are counter-based, so pass 2 can recompute them)
//   scan_*_kernel         exclusive prefix sum over the owned cells in cell-id order -> first record of every cell
//   mutate_emit_kernel    recomputes the draws, ranks the events of a cell by slot (warp ballots + a per-tile running
//                         base in shared memory), writes each record at its final position and subtracts the mutants
// ---------------------------------------------------------------------------------------------
struct MutRec { uint32_t cell, slot, count; };

// born counts of this thread's 4 cells of species row s (0 where nothing happens)
__device__ __forceinline__ void mutate_born(const Layout& L, const int4 v, double m, uint32_t xq, uint32_t yg, uint32_t s,
                                            unsigned long long t, uint32_t seed, int (&born)[4]) {
  born[0] = born[1] = born[2] = born[3] = 0;
  if ((v.x | v.y | v.z | v.w) == 0 || !(m > 0.0)) return;   // extinct here / ROUND(x <= 0) = 0 without a draw
  const uint4 u = draw4(xq, yg, s, t, seed, 1u);
  const int n[4] = {v.x, v.y, v.z, v.w};
  const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (n[k] > 0 && xq * 4 + k < L.nx) born[k] = round_stoch(__dmul_rn(m, u2d((uint32_t)n[k])), uu[k]);
}

// pass 1: counts[yl*nx + x] = number of species slots of that cell with born > 0.  One CTA per tile at a time.
__global__ void __launch_bounds__(256) mutate_count_kernel(const Layout L, const double* __restrict__ ms, uint32_t seed,
                                                           unsigned long long t, uint32_t* __restrict__ counts) {
  __shared__ uint32_t cnt[32];
  const uint32_t ntiles = L.ntx * L.ny, qpr = L.tw / 4, per = L.nsp * qpr;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    if (threadIdx.x < 32) cnt[threadIdx.x] = 0;
    __syncthreads();
    const int4* tp = reinterpret_cast<const int4*>(L.base + (size_t)tile * L.tile_stride);
    const uint32_t yl = tile / L.ntx, tx = tile - yl * L.ntx, yg = L.y0 + yl;
    uint32_t c[4] = {0, 0, 0, 0};
    const uint32_t qd = threadIdx.x % qpr;      // blockDim is a multiple of qpr: a thread always sees the same 4 cell columns
    for (uint32_t i = threadIdx.x; i < per; i += blockDim.x) {
      const uint32_t s = i / qpr;
      int born[4];
      mutate_born(L, tp[i], __ldg(ms + s), tx * qpr + qd, yg, s, t, seed, born);
#pragma unroll
      for (int k = 0; k < 4; ++k) c[k] += born[k] > 0;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (c[k]) atomicAdd(&cnt[4 * qd + k], c[k]);   // shared-memory atomics only
    __syncthreads();
    const uint32_t x = tx * L.tw + threadIdx.x;
    if (threadIdx.x < L.tw && x < L.nx) counts[(size_t)yl * L.nx + x] = cnt[threadIdx.x];
    __syncthreads();
  }
}

// exclusive scan of n counts, three launches: per-block scans + block totals; scan of the totals (one CTA); the block
// prefix is added by the reader (offset(i) = blocksum[i / 1024] + local[i]).  Hand-written: n <= 2^24 cells here.
constexpr uint32_t kScanBlock = 1024;
__device__ __forceinline__ uint32_t block_exclusive_scan_1024(uint32_t v, uint32_t* warp_tot /*[32] shared*/, uint32_t& total) {
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(FULL, inc, o);
    if (lane >= (uint32_t)o) inc += up;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = warp_tot[lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(FULL, winc, o);
      if (lane >= (uint32_t)o) winc += up;
    }
    warp_tot[lane] = winc - w;            // exclusive prefix of the warp totals
    if (lane == 31) warp_tot[32] = winc;  // block total
  }
  __syncthreads();
  total = warp_tot[32];
  const uint32_t r = warp_tot[warp] + inc - v;
  __syncthreads();
  return r;
}
__global__ void __launch_bounds__(kScanBlock) scan_local_kernel(uint32_t* __restrict__ data, size_t n, uint32_t* __restrict__ blocksum) {
  __shared__ uint32_t wt[33];
  const size_t i = (size_t)blockIdx.x * kScanBlock + threadIdx.x;
  const uint32_t v = i < n ? data[i] : 0u;
  uint32_t total;
  const uint32_t ex = block_exclusive_scan_1024(v, wt, total);
  if (i < n) data[i] = ex;
  if (threadIdx.x == 0) blocksum[blockIdx.x] = total;
}
// exclusive scan of the block totals in place (nb <= 2^24 / 1024 = 16384, a few trips of one CTA); total -> *grand
__global__ void __launch_bounds__(kScanBlock) scan_tops_kernel(uint32_t* __restrict__ blocksum, uint32_t nb, unsigned long long* grand) {
  __shared__ uint32_t wt[33];
  unsigned long long carry = 0;
  for (uint32_t base = 0; base < nb; base += kScanBlock) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < nb ? blocksum[i] : 0u;
    uint32_t total;
    const uint32_t ex = block_exclusive_scan_1024(v, wt, total);
    if (i < nb) blocksum[i] = (uint32_t)carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) *grand = carry;
}

// pass 2: every event written at its final position, in (cell, slot) order; densities reduced by the mutants.
// A CTA walks a tile in chunks of blockDim/qpr species rows.  Within a chunk the rank of an event among the events of
// the same cell = events in lower warps (shared-memory table) + events in lower lanes of the same quad (ballot).
__global__ void __launch_bounds__(256) mutate_emit_kernel(const Layout L, const double* __restrict__ ms, uint32_t seed,
                                                          unsigned long long t, const uint32_t* __restrict__ offs,
                                                          const uint32_t* __restrict__ blocksum, MutRec* __restrict__ rec) {
  __shared__ uint32_t base[32];          // per cell of the tile: records already written
  __shared__ uint32_t wcnt[8][32];       // per warp, per cell: events of the current chunk
  const uint32_t ntiles = L.ntx * L.ny, qpr = L.tw / 4, per = L.nsp * qpr;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const uint32_t qd = threadIdx.x % qpr;
  // lanes of this warp that hold the same 4 cell columns and a lower species row
  uint32_t same_quad = 0;
  for (uint32_t l = 0; l < 32; ++l)
    if (l % qpr == lane % qpr) same_quad |= 1u << l;
  const uint32_t lower = same_quad & ((1u << lane) - 1u);
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int4* tp = reinterpret_cast<int4*>(L.base + (size_t)tile * L.tile_stride);
    const uint32_t yl = tile / L.ntx, tx = tile - yl * L.ntx, yg = L.y0 + yl;
    if (threadIdx.x < 32) {
      const uint32_t x = tx * L.tw + threadIdx.x;
      uint32_t b = 0;
      if (threadIdx.x < L.tw && x < L.nx) {
        const size_t c = (size_t)yl * L.nx + x;
        b = blocksum[c / kScanBlock] + offs[c];
      }
      base[threadIdx.x] = b;
    }
    __syncthreads();
    for (uint32_t i0 = 0; i0 < per; i0 += blockDim.x) {     // uniform trip count: every thread reaches the barriers
      const uint32_t i = i0 + threadIdx.x;
      const uint32_t s = i / qpr;
      int born[4] = {0, 0, 0, 0};
      int4 v = make_int4(0, 0, 0, 0);
      if (i < per) {
        ECO_CHK(tp + i, 16, 41u);
        v = tp[i];
        mutate_born(L, v, __ldg(ms + s), tx * qpr + qd, yg, s, t, seed, born);
      }
      uint32_t rank[4];
      bool any = false;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t bal = __ballot_sync(FULL, born[k] > 0);
        rank[k] = __popc(bal & lower);
        const uint32_t tot = __popc(bal & same_quad);
        if (lane < qpr) wcnt[warp][4 * lane + k] = tot;     // lane < qpr: one writer per (quad, k); lane % qpr == lane
        any |= born[k] > 0;
      }
      __syncthreads();
      if (any) {
        int n[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (born[k] > 0) {
            uint32_t pos = base[4 * qd + k] + rank[k];
            for (uint32_t w = 0; w < warp; ++w) pos += wcnt[w][4 * qd + k];
            ECO_CHK(rec + pos, sizeof(MutRec), 43u);
            rec[pos] = MutRec{yg * L.nx + tx * L.tw + 4 * qd + k, s, (uint32_t)born[k]};
            n[k] -= born[k];
          }
        ECO_CHK(tp + i, 16, 42u);
        tp[i] = make_int4(n[0], n[1], n[2], n[3]);
      }
      __syncthreads();
      if (threadIdx.x < L.tw) {
        uint32_t add = 0;
        for (uint32_t w = 0; w < nwarps; ++w) add += wcnt[w][threadIdx.x];
        base[threadIdx.x] += add;
      }
      __syncthreads();
    }
  }
}

// new species rows: zero everywhere, then 1 in the birth cell (reference :272-275)
__global__ void zero_rows_kernel(const Layout L, uint32_t ntiles_all, int32_t* buf, uint32_t s0, uint32_t s1) {
  const uint32_t per_tile = (s1 - s0) * L.tw;
  const unsigned long long n = (unsigned long long)ntiles_all * per_tile;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t tile = (uint32_t)(i / per_tile), o = (uint32_t)(i % per_tile);
    ECO_CHK(&buf[(size_t)tile * L.tile_stride + (size_t)s0 * L.tw + o], 4, 40u);
    buf[(size_t)tile * L.tile_stride + (size_t)s0 * L.tw + o] = 0;
  }
}
__global__ void birth_kernel(const Layout L, uint32_t s0, const uint32_t* cells, uint32_t m) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint32_t c = cells[i];
  const uint32_t yg = c / L.nx, x = c - yg * L.nx;
  if (yg < L.y0 || yg >= L.y0 + L.ny) return;
  ECO_CHK(&L.base[cell_addr(L, yg - L.y0, x, s0 + i)], 4, 33u);
  L.base[cell_addr(L, yg - L.y0, x, s0 + i)] = 1;
}

// ---------------------------------------------------------------------------------------------
// observables
// ---------------------------------------------------------------------------------------------
// living species per owned cell (EcolabPoint::nsp, models/ecolab_model.cc:114-122); one CTA per tile, 16-byte loads
__global__ void __launch_bounds__(256) nsp_cell_kernel(const Layout L, uint32_t* out) {
  __shared__ uint32_t cnt[32];
  const uint32_t ntiles = L.ntx * L.ny, qpr = L.tw / 4, per = L.nsp * qpr;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    if (threadIdx.x < 32) cnt[threadIdx.x] = 0;
    __syncthreads();
    const int4* t = reinterpret_cast<const int4*>(L.base + (size_t)tile * L.tile_stride);
    uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;   // blockDim is a multiple of qpr: a thread always sees the same 4 cell columns
#pragma unroll 4
    for (uint32_t i = threadIdx.x; i < per; i += blockDim.x) {
      const int4 v = __ldcs(t + i);
      c0 += v.x != 0; c1 += v.y != 0; c2 += v.z != 0; c3 += v.w != 0;
    }
    const uint32_t q = threadIdx.x % qpr;
    atomicAdd(&cnt[4 * q], c0); atomicAdd(&cnt[4 * q + 1], c1); atomicAdd(&cnt[4 * q + 2], c2); atomicAdd(&cnt[4 * q + 3], c3);
    __syncthreads();
    const uint32_t yl = tile / L.ntx, tx = tile - yl * L.ntx, x = tx * L.tw + threadIdx.x;
    if (threadIdx.x < L.tw && x < L.nx) out[(size_t)yl * L.nx + x] = cnt[threadIdx.x];
    __syncthreads();
  }
}

// per-species totals over owned cells: CTA = (chunk of species rows, range of tiles); out must be zeroed
__global__ void __launch_bounds__(1024) totals_kernel(const Layout L, uint32_t tiles_per_cta, unsigned long long* out) {
  const uint32_t ch = blockDim.x / L.tw;                     // species rows per chunk
  const uint32_t nchunks = (L.nsp + ch - 1) / ch;
  const uint32_t chunk = blockIdx.x % nchunks, part = blockIdx.x / nchunks;
  const uint32_t s = chunk * ch + threadIdx.x / L.tw, ntiles = L.ntx * L.ny;
  const uint32_t t0 = part * tiles_per_cta, t1 = min(ntiles, t0 + tiles_per_cta);
  long long acc = 0;
  if (s < L.nsp)
    for (uint32_t tile = t0; tile < t1; ++tile) acc += L.base[(size_t)tile * L.tile_stride + (size_t)s * L.tw + threadIdx.x % L.tw];
  for (uint32_t o = L.tw / 2; o; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
  if (s < L.nsp && threadIdx.x % L.tw == 0 && acc) atomicAdd(out + s, (unsigned long long)acc);
}

// ---------------------------------------------------------------------------------------------
// layout conversion: cell-major [cell][nsp] <-> blocked; capacity / tile-width changes; fills
// ---------------------------------------------------------------------------------------------
template <bool TO_BLOCKED>
__global__ void __launch_bounds__(1024) transpose_kernel(const Layout L, int32_t* cellmajor) {
  __shared__ int32_t sm[32][33];
  const uint32_t tx_ = threadIdx.x & 31u, ty_ = threadIdx.x >> 5;
  const uint32_t nxb = (L.nx + 31) / 32, nsb = (L.nsp + 31) / 32;
  const unsigned long long nwork = (unsigned long long)nxb * L.ny * nsb;
  for (unsigned long long w = blockIdx.x; w < nwork; w += gridDim.x) {
    const uint32_t sb = (uint32_t)(w % nsb);
    const uint32_t xb = (uint32_t)((w / nsb) % nxb), yl = (uint32_t)(w / ((unsigned long long)nsb * nxb));
    if (TO_BLOCKED) {
      const uint32_t x = xb * 32u + ty_, s = sb * 32u + tx_;     // read coalesced along species
      sm[ty_][tx_] = (x < L.nx && s < L.nsp) ? cellmajor[((size_t)yl * L.nx + x) * L.nsp + s] : 0;
      __syncthreads();
      const uint32_t s2 = sb * 32u + ty_, x2 = xb * 32u + tx_;   // write coalesced along cells
      if (s2 < L.nsp && x2 < L.ntx * L.tw) { ECO_CHK(&L.base[cell_addr(L, yl, x2, s2)], 4, 34u); L.base[cell_addr(L, yl, x2, s2)] = (x2 < L.nx) ? sm[tx_][ty_] : 0; }
      __syncthreads();
    } else {
      const uint32_t s2 = sb * 32u + ty_, x2 = xb * 32u + tx_;
      if (s2 < L.nsp && x2 < L.nx) ECO_CHK(&L.base[cell_addr(L, yl, x2, s2)], 4, 35u);
      sm[ty_][tx_] = (s2 < L.nsp && x2 < L.nx) ? L.base[cell_addr(L, yl, x2, s2)] : 0;
      __syncthreads();
      const uint32_t x = xb * 32u + ty_, s = sb * 32u + tx_;
      if (x < L.nx && s < L.nsp) cellmajor[((size_t)yl * L.nx + x) * L.nsp + s] = sm[tx_][ty_];
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------------------
// beta: CSR -> the blocked-ELL record stream of k_generate.cuh.  Block k = slots [boff[k], boff[k+1]) of R 16-byte records:
// slot 0 holds {diag, repro_rate} of the block's R species rows, slot 1 + e the e-th term of each row
// {val, byte offset of the term's species row inside a tile, high word of c = -(val * 2^52)}; everything else is the padding
// record {0.0, offset 0, -0.0} (a term that adds exactly nothing).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ell_build_kernel(const uint32_t* __restrict__ boff, uint32_t nblocks, uint32_t R, uint32_t nsp, uint32_t tw,
                                                        const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ col,
                                                        const double* __restrict__ val, const double* __restrict__ diag,
                                                        const double* __restrict__ rr, uint4* __restrict__ stream) {
  for (uint32_t k = blockIdx.x; k < nblocks; k += gridDim.x) {
    const uint32_t s0 = boff[k], nrec = (boff[k + 1] - s0) * R;
    for (uint32_t i = threadIdx.x; i < nrec; i += blockDim.x) {
      const uint32_t slot = i / R, row = k * R + i % R;
      uint4 rec = make_uint4(0u, 0u, 0u, 0x80000000u);
      if (row < nsp) {
        if (slot == 0) {
          const double d = diag[row], q = rr[row];
          rec = make_uint4((uint32_t)__double2loint(d), (uint32_t)__double2hiint(d), (uint32_t)__double2loint(q), (uint32_t)__double2hiint(q));
        } else {
          const uint32_t e = rowptr[row] + slot - 1u;
          if (e < rowptr[row + 1]) {
            const double v = val[e], c = -__dmul_rn(v, 4503599627370496.0);
            rec = make_uint4((uint32_t)__double2loint(v), (uint32_t)__double2hiint(v), col[e] * tw * 4u, (uint32_t)__double2hiint(c));
          }
        }
      }
      stream[(size_t)s0 * R + i] = rec;
    }
  }
}

// copy every (row incl. ghosts, cell, species < nrows) from one blocked layout to another (different cap / tw);
// the destination has been zero-filled
__global__ void relayout_kernel(const Layout S, const Layout D, uint32_t nrows_y, uint32_t nrows_s) {
  const uint32_t xw = min(S.ntx * S.tw, D.ntx * D.tw);
  const unsigned long long n = (unsigned long long)nrows_y * nrows_s * xw;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t x = (uint32_t)(i % xw);
    const uint32_t s = (uint32_t)((i / xw) % nrows_s), y = (uint32_t)(i / ((unsigned long long)xw * nrows_s));
    ECO_CHK(&D.base[cell_addr(D, y, x, s)], 4, 36u);
    ECO_CHK(&S.base[cell_addr(S, y, x, s)], 4, 37u);
    D.base[cell_addr(D, y, x, s)] = S.base[cell_addr(S, y, x, s)];
  }
}

__global__ void fill_kernel(const Layout L, int32_t v) {
  const uint32_t per_tile = L.nsp * L.tw, ntiles = L.ntx * L.ny;
  const unsigned long long n = (unsigned long long)ntiles * per_tile;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t tile = (uint32_t)(i / per_tile), o = (uint32_t)(i % per_tile);
    const uint32_t tx = tile % L.ntx;
    ECO_CHK(&L.base[(size_t)tile * L.tile_stride + o], 4, 38u);
    L.base[(size_t)tile * L.tile_stride + o] = (tx * L.tw + o % L.tw < L.nx) ? v : 0;
  }
}

// synthetic initial state: n = lo + draw % span with the stage-3 Philox stream at t = 0
__global__ void __launch_bounds__(256) fill_random_kernel(const Layout L, uint32_t seed, int32_t lo, uint32_t span) {
  const uint32_t qpr = L.tw / 4, per = L.nsp * qpr, ntiles = L.ntx * L.ny;
  const unsigned long long n = (unsigned long long)ntiles * per;
  for (unsigned long long w = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; w < n;
       w += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t tile = (uint32_t)(w / per), i = (uint32_t)(w % per);
    const uint32_t s = i / qpr, qd = i - s * qpr;
    const uint32_t yl = tile / L.ntx, tx = tile - yl * L.ntx, xq = tx * qpr + qd;
    const uint4 u = draw4(xq, L.y0 + yl, s, 0ull, seed, 3u);
    int4 v;
    v.x = xq * 4 + 0 < L.nx ? lo + (int32_t)(u.x % span) : 0;
    v.y = xq * 4 + 1 < L.nx ? lo + (int32_t)(u.y % span) : 0;
    v.z = xq * 4 + 2 < L.nx ? lo + (int32_t)(u.z % span) : 0;
    v.w = xq * 4 + 3 < L.nx ? lo + (int32_t)(u.w % span) : 0;
    ECO_CHK(reinterpret_cast<int4*>(L.base + (size_t)tile * L.tile_stride) + i, 16, 39u);
    reinterpret_cast<int4*>(L.base + (size_t)tile * L.tile_stride)[i] = v;
  }
}

}  // namespace eco
