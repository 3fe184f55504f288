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
          v[u] = *reinterpret_cast<const int4*>(t + (size_t)sp * tw);
          if (__ldg(keep + sp)) dst[u] = __ldg(map + sp);
        }
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (dst[u] != 0xffffffffu) *reinterpret_cast<int4*>(t + (size_t)dst[u] * tw) = v[u];
      __syncwarp();
    }
    for (uint32_t sp = nkeep + rr; sp < nsp_old; sp += rows) *reinterpret_cast<int4*>(t + (size_t)sp * tw) = make_int4(0, 0, 0, 0);
  }
}

// ---------------------------------------------------------------------------------------------
// mutate, cell side (reference EcolabPoint::mutate, models/ecolab_model.cc:278-314):
//   k = ROUND(ms[s] * n[c][s]) with the stage-1 Philox stream; n -= k; record (cell, slot, k>0)
// ---------------------------------------------------------------------------------------------
struct MutRec { uint32_t cell, slot, count; };

__global__ void __launch_bounds__(256) mutate_scan_kernel(const Layout L, const double* ms, uint32_t seed,
                                                          unsigned long long t, MutRec* rec, unsigned long long rec_cap,
                                                          unsigned long long* count) {
  const uint32_t ntiles = L.ntx * L.ny, qpr = L.tw / 4, per = L.nsp * qpr;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int4* tp = reinterpret_cast<int4*>(L.base + (size_t)tile * L.tile_stride);
    const uint32_t yl = tile / L.ntx, tx = tile - yl * L.ntx, yg = L.y0 + yl;
    for (uint32_t i = threadIdx.x; i < per; i += blockDim.x) {
      const uint32_t s = i / qpr, qd = i - s * qpr;
      int4 v = tp[i];
      if ((v.x | v.y | v.z | v.w) == 0) continue;   // densities are non-negative: all four extinct here
      const double m = __ldg(ms + s);
      if (!(m > 0.0)) continue;                     // ROUND(x <= 0) = 0 without a draw
      const uint32_t xq = tx * qpr + qd;
      const uint4 u = draw4(xq, yg, s, t, seed, 1u);
      int n[4] = {v.x, v.y, v.z, v.w};
      const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
      bool any = false;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t x = xq * 4 + k;
        if (n[k] > 0 && x < L.nx) {
          const int born = round_stoch(__dmul_rn(m, u2d((uint32_t)n[k])), uu[k]);
          if (born > 0) {
            const unsigned long long slot = atomicAdd(count, 1ull);
            if (slot < rec_cap) { rec[slot] = MutRec{yg * L.nx + x, s, (uint32_t)born}; n[k] -= born; any = true; }
          }
        }
      }
      if (any) tp[i] = make_int4(n[0], n[1], n[2], n[3]);
    }
  }
}

// undo a scan whose record buffer overflowed (the scan is then repeated with a larger buffer)
__global__ void mutate_undo_kernel(const Layout L, const MutRec* rec, uint32_t nrec) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nrec) return;
  const MutRec r = rec[i];
  const uint32_t yg = r.cell / L.nx, x = r.cell - yg * L.nx;
  L.base[cell_addr(L, yg - L.y0, x, r.slot)] += (int)r.count;
}

// new species rows: zero everywhere, then 1 in the birth cell (reference :272-275)
__global__ void zero_rows_kernel(const Layout L, uint32_t ntiles_all, int32_t* buf, uint32_t s0, uint32_t s1) {
  const uint32_t per_tile = (s1 - s0) * L.tw;
  const unsigned long long n = (unsigned long long)ntiles_all * per_tile;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t tile = (uint32_t)(i / per_tile), o = (uint32_t)(i % per_tile);
    buf[(size_t)tile * L.tile_stride + (size_t)s0 * L.tw + o] = 0;
  }
}
__global__ void birth_kernel(const Layout L, uint32_t s0, const uint32_t* cells, uint32_t m) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint32_t c = cells[i];
  const uint32_t yg = c / L.nx, x = c - yg * L.nx;
  if (yg < L.y0 || yg >= L.y0 + L.ny) return;
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
      if (s2 < L.nsp && x2 < L.ntx * L.tw) L.base[cell_addr(L, yl, x2, s2)] = (x2 < L.nx) ? sm[tx_][ty_] : 0;
      __syncthreads();
    } else {
      const uint32_t s2 = sb * 32u + ty_, x2 = xb * 32u + tx_;
      sm[ty_][tx_] = (s2 < L.nsp && x2 < L.nx) ? L.base[cell_addr(L, yl, x2, s2)] : 0;
      __syncthreads();
      const uint32_t x = xb * 32u + ty_, s = sb * 32u + tx_;
      if (x < L.nx && s < L.nsp) cellmajor[((size_t)yl * L.nx + x) * L.nsp + s] = sm[tx_][ty_];
      __syncthreads();
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
    reinterpret_cast<int4*>(L.base + (size_t)tile * L.tile_stride)[i] = v;
  }
}

}  // namespace eco
