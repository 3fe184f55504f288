// k_generate.cuh -- the generalised Lotka-Volterra step with stochastic rounding
// (reference EcolabPoint::generate, models/ecolab_model.cc:90-112; ROUND :57-66).
//
// beta on the device is a blocked-ELL *stream* matched to the lane map.  Species rows are processed
// ROWS = 128/TW at a time (one warp-op: ROWS rows x TW cells, a lane owns 4 consecutive cells of one row).
// Block b occupies the 16-byte records  stream[boff[b]*ROWS ... boff[b+1]*ROWS):
//     slot 0      : ROWS headers  { double diag ; double repro_rate }
//     slot 1..W_b : ROWS terms    { double val ; uint32 byte offset of species row `col` inside a tile ; pad }
// where W_b = longest row of the block (shorter rows are padded with val = 0, offset 0).
// A warp-op therefore reads its whole beta block through ONE base pointer with immediate offsets.
// Terms of a row keep the COO order (reference oDiagIdx, :718-723) so the FP64 summation order is the reference's:
//     ir = diag*n ; ir = ir + val_j*n[col_j] (j ascending) ; x = n + n*(r + ir), every operation individually rounded.
//
#pragma once
#include "k_common.cuh"

namespace eco {

struct GenParams {
  Layout L;
  uint32_t nblocks;          // ceil(nsp / ROWS)
  const uint32_t* boff;      // [nblocks+1] slot offsets of the blocks inside `stream`
  const uint4* stream;       // beta stream, see above
  uint32_t seed, niter;
  unsigned long long t0;
  uint32_t nstages;          // pipeline kernel: shared-memory tile buffers
  uint32_t stream_recs;      // records in `stream` (BSMEM kernels copy it into shared memory)
  uint32_t rk[20];           // Philox round keys {k0_r, k1_r}, r = 0..9, for (seed, t0 >> 32): host-computed
};

template <bool BSMEM>
__device__ __forceinline__ uint4 ldrec(const uint4* p) {
  if (BSMEM) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
    return v;
  }
  return __ldg(p);
}

// Philox4x32-10 with the key schedule taken from kernel parameters (constant bank operands, no key arithmetic)
__device__ __forceinline__ uint4 philox_rk(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const uint32_t (&rk)[20]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned long long p0 = (unsigned long long)0xD2511F53u * c0;
    const unsigned long long p1 = (unsigned long long)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ rk[2 * r];
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ rk[2 * r + 1];
    c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
  }
  return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ double dbl(uint32_t lo, uint32_t hi) { return __hiloint2double((int)hi, (int)lo); }

// One warp-op: the four pre-rounding values x = n + n*(r + ir) of this lane's cells of species row `row`.
// `lp` = shared-memory tile + 4*quad (ints); `own` = the lane's own densities; `blk` = this lane's header record of
// block b (global memory, or the shared-memory copy when BSMEM); W = terms per row.
template <int TW, bool BSMEM>
__device__ __forceinline__ void gen_x(const int32_t* __restrict__ lp, const int4 own, const uint4* __restrict__ blk,
                                      uint32_t W, double x[4]) {
  constexpr int R = LaneMap<TW>::ROWS;
  const uint4 h = ldrec<BSMEM>(blk);
  const double diag = dbl(h.x, h.y), rr = dbl(h.z, h.w);
  const double d0 = (double)(uint32_t)own.x, d1 = (double)(uint32_t)own.y, d2 = (double)(uint32_t)own.z,
               d3 = (double)(uint32_t)own.w;
  double ir0 = __dmul_rn(diag, d0), ir1 = __dmul_rn(diag, d1), ir2 = __dmul_rn(diag, d2), ir3 = __dmul_rn(diag, d3);
  const uint4* e = blk + R;
#pragma unroll 4
  for (uint32_t j = 0; j < W; ++j, e += R) {
    const uint4 w = ldrec<BSMEM>(e);
    const double v = dbl(w.x, w.y);
    const int4 g = *reinterpret_cast<const int4*>(reinterpret_cast<const char*>(lp) + w.z);
    // int -> double on the conversion (XU) pipe, which is otherwise idle; multiply and accumulate stay two
    // individually rounded FP64 operations (no contraction: -fmad=false)
    ir0 = __dadd_rn(ir0, __dmul_rn(v, (double)(uint32_t)g.x));
    ir1 = __dadd_rn(ir1, __dmul_rn(v, (double)(uint32_t)g.y));
    ir2 = __dadd_rn(ir2, __dmul_rn(v, (double)(uint32_t)g.z));
    ir3 = __dadd_rn(ir3, __dmul_rn(v, (double)(uint32_t)g.w));
  }
  x[0] = __dadd_rn(d0, __dmul_rn(d0, __dadd_rn(rr, ir0)));
  x[1] = __dadd_rn(d1, __dmul_rn(d1, __dadd_rn(rr, ir1)));
  x[2] = __dadd_rn(d2, __dmul_rn(d2, __dadd_rn(rr, ir2)));
  x[3] = __dadd_rn(d3, __dmul_rn(d3, __dadd_rn(rr, ir3)));
}

// ROUND of the lane's four values.  Fast path: 0 <= x < 2^31-256 for all four (decided on the high words with
// integer compares, one warp-uniform branch); otherwise the fully guarded round_stoch.
__device__ __forceinline__ int4 round4(const double x[4], const uint4 u) {
  const uint32_t h0 = (uint32_t)__double2hiint(x[0]), h1 = (uint32_t)__double2hiint(x[1]),
                 h2 = (uint32_t)__double2hiint(x[2]), h3 = (uint32_t)__double2hiint(x[3]);
  const uint32_t hm = max(max(h0, h1), max(h2, h3));     // negative x has the sign bit set -> huge as unsigned
  int4 o;
  if (__any_sync(FULL, hm >= 0x41DFFFFFu)) {
    o.x = round_stoch(x[0], u.x); o.y = round_stoch(x[1], u.y);
    o.z = round_stoch(x[2], u.z); o.w = round_stoch(x[3], u.w);
  } else {
    o.x = round_stoch_inrange(x[0], u.x); o.y = round_stoch_inrange(x[1], u.y);
    o.z = round_stoch_inrange(x[2], u.z); o.w = round_stoch_inrange(x[3], u.w);
  }
  return o;
}

__device__ __forceinline__ void load_tile_bulk(int32_t* dst_smem, const int32_t* src, uint32_t tile_bytes, uint64_t* bar) {
  mbar_expect_tx(bar, tile_bytes);
  for (uint32_t off = 0; off < tile_bytes; off += 32768u) {   // modest requests, one barrier
    const uint32_t n = min(32768u, tile_bytes - off);
    bulk_g2s(reinterpret_cast<unsigned char*>(dst_smem) + off, reinterpret_cast<const unsigned char*>(src) + off, n, bar);
  }
}

// ---------------------------------------------------------------------------------------------------------
// generate(1): persistent producer/consumer pipeline, one CTA per SM.
//   warp NW        : producer -- one lane streams the CTA's tiles (all species of TW cells, one contiguous
//                    block in HBM) into a ring of `nstages` shared-memory buffers with bulk (TMA-unit) copies,
//                    completion on per-stage "full" mbarriers; it re-fills a buffer as soon as all consumer
//                    warps have arrived on its "empty" mbarrier.
//   warps 0..NW-1  : consumers -- warp-ops (ROWS species rows x TW cells) are dealt round-robin over the warps
//                    across tile boundaries, so the load stays balanced when nblocks is not a multiple of NW and
//                    no CTA-wide barrier exists: a warp that is done with a tile moves on to the next one.
// Results go straight from registers to HBM as 16-byte vectors (a warp-op writes 512 contiguous bytes), in place.
// ---------------------------------------------------------------------------------------------------------
template <int TW, int NW, bool BSMEM>
__global__ void __launch_bounds__((NW + 1) * 32, 1) generate_pipe_kernel(const GenParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int R = LaneMap<TW>::ROWS;
  const uint32_t nblocks = p.nblocks, S = p.nstages;
  const uint32_t tile_words = nblocks * R * TW, tile_bytes = tile_words * 4u;
  int32_t* tiles = reinterpret_cast<int32_t*>(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)S * tile_bytes);
  uint64_t* empty = full + S;
  uint32_t* s_boff = reinterpret_cast<uint32_t*>(empty + S);
  uint4* s_stream = reinterpret_cast<uint4*>(smem_raw + (((size_t)S * tile_bytes + 16 * S + 4 * (nblocks + 1)) + 15) / 16 * 16);

  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t i = threadIdx.x; i <= nblocks; i += (NW + 1) * 32) s_boff[i] = p.boff[i];
  if (BSMEM)
    for (uint32_t i = threadIdx.x; i < p.stream_recs; i += (NW + 1) * 32) s_stream[i] = __ldg(p.stream + i);
  const uint4* strm = BSMEM ? s_stream : p.stream;
  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < S; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, NW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const uint32_t ntiles = p.L.ntx * p.L.ny;
  if (warp == NW) {
    if (lane == 0) {
      uint32_t stage = 0, round = 0;
      for (uint32_t tile_id = blockIdx.x; tile_id < ntiles; tile_id += gridDim.x) {
        if (round) mbar_wait(empty + stage, (round - 1u) & 1u);
        load_tile_bulk(tiles + (size_t)stage * tile_words, p.L.base + (size_t)tile_id * p.L.tile_stride, tile_bytes, full + stage);
        if (++stage == S) { stage = 0; ++round; }
      }
    }
    return;
  }

  const uint32_t rr = LaneMap<TW>::row(lane), q = LaneMap<TW>::quad(lane);
  const uint32_t nbm = nblocks % NW;
  uint32_t first = warp, stage = 0, round = 0;
  for (uint32_t tile_id = blockIdx.x; tile_id < ntiles; tile_id += gridDim.x) {
    mbar_wait(full + stage, round & 1u);
    const int32_t* lp = tiles + (size_t)stage * tile_words + 4 * q;
    int32_t* gl = p.L.base + (size_t)tile_id * p.L.tile_stride + 4 * q;
    const uint32_t yl = tile_id / p.L.ntx, tx = tile_id - yl * p.L.ntx;
    const uint32_t xq = tx * (TW / 4) + q, yg = p.L.y0 + yl;
    for (uint32_t b = first; b < nblocks; b += NW) {
      const uint32_t row = b * R + rr;
      const int4 own = *reinterpret_cast<const int4*>(lp + row * TW);
      // n = 0 stays 0 (x = 0 + 0*(...) = 0, ROUND(0) = 0): a warp-op whose 128 densities are all extinct is skipped
      if (!__any_sync(FULL, (own.x | own.y | own.z | own.w) != 0)) continue;
      const uint32_t s0 = s_boff[b], s1 = s_boff[b + 1];
      double x[4];
#ifndef ECO_ABLATE
      gen_x<TW, BSMEM>(lp, own, strm + (size_t)s0 * R + rr, s1 - s0 - 1u, x);
      const int4 o = round4(x, philox_rk(xq, yg, row, (uint32_t)p.t0, p.rk));
      __stcs(reinterpret_cast<int4*>(gl + row * TW), o);
#else   // cost-attribution builds (tools/ablate.sh): results are wrong on purpose, never shipped
      gen_x<TW, BSMEM>(lp, own, strm + (size_t)s0 * R + rr, (ECO_ABLATE & 2) ? 0u : s1 - s0 - 1u, x);
      const uint4 u = (ECO_ABLATE & 1) ? make_uint4(row, row * 3u, xq, yg) : philox_rk(xq, yg, row, (uint32_t)p.t0, p.rk);
      int4 o;
      if (ECO_ABLATE & 4) o = make_int4(__double2loint(x[0]) ^ u.x, __double2loint(x[1]) ^ u.y, __double2loint(x[2]) ^ u.z, __double2loint(x[3]) ^ u.w);
      else o = round4(x, u);
      if (!(ECO_ABLATE & 8) || o.x == 0x7fffffff) __stcs(reinterpret_cast<int4*>(gl + row * TW), o);
#endif
    }
    first += NW - nbm;
    if (first >= NW) first -= NW;
    __syncwarp();
    if (lane == 0) mbar_arrive(empty + stage);
    if (++stage == S) { stage = 0; ++round; }
  }
}

// ---------------------------------------------------------------------------------------------------------
// generate(niter > 1): per tile, `niter` Jacobi steps between two shared-memory buffers (HBM traffic amortised
// over niter -- the scripts' cadence is generate(100), models/spatial_ecolab.py:67), CTA-wide barrier per step.
// ---------------------------------------------------------------------------------------------------------
template <int TW, int THREADS, bool BSMEM>
__global__ void __launch_bounds__(THREADS, 1) generate_multi_kernel(const GenParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int R = LaneMap<TW>::ROWS;
  const uint32_t nblocks = p.nblocks;
  const uint32_t tile_words = nblocks * R * TW, tile_bytes = tile_words * 4u;
  int32_t* buf0 = reinterpret_cast<int32_t*>(smem_raw);
  int32_t* buf1 = buf0 + tile_words;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + (size_t)tile_bytes * 2);
  uint32_t* s_boff = reinterpret_cast<uint32_t*>(bar + 1);
  uint4* s_stream = reinterpret_cast<uint4*>(smem_raw + (((size_t)tile_bytes * 2 + 8 + 4 * (nblocks + 1)) + 15) / 16 * 16);

  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = THREADS / 32;
  const uint32_t rr = LaneMap<TW>::row(lane), q = LaneMap<TW>::quad(lane);
  for (uint32_t i = threadIdx.x; i <= nblocks; i += THREADS) s_boff[i] = p.boff[i];
  if (BSMEM)
    for (uint32_t i = threadIdx.x; i < p.stream_recs; i += THREADS) s_stream[i] = __ldg(p.stream + i);
  const uint4* strm = BSMEM ? s_stream : p.stream;
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
      load_tile_bulk(buf0, gt, tile_bytes, bar);
      const uint32_t nxt = tile_id + gridDim.x;
      if (nxt < ntiles) bulk_prefetch_l2(p.L.base + (size_t)nxt * p.L.tile_stride, tile_bytes);
    }
    mbar_wait(bar, phase);
    phase ^= 1u;
    const uint32_t yl = tile_id / p.L.ntx, tx = tile_id - yl * p.L.ntx;
    const uint32_t xq = tx * (TW / 4) + q, yg = p.L.y0 + yl;
    int32_t* src = buf0;
    int32_t* dst = buf1;
    for (uint32_t step = 0; step < p.niter; ++step) {
      const int32_t* lp = src + 4 * q;
      for (uint32_t b = warp; b < nblocks; b += nwarps) {
        const uint32_t row = b * R + rr;
        const int4 own = *reinterpret_cast<const int4*>(lp + row * TW);
        int4 o = make_int4(0, 0, 0, 0);
        if (__any_sync(FULL, (own.x | own.y | own.z | own.w) != 0)) {
          const uint32_t s0 = s_boff[b], s1 = s_boff[b + 1];
          double x[4];
          gen_x<TW, BSMEM>(lp, own, strm + (size_t)s0 * R + rr, s1 - s0 - 1u, x);
          o = round4(x, philox_rk(xq, yg, row, (uint32_t)p.t0 + step, p.rk));
        }
        *reinterpret_cast<int4*>(dst + row * TW + 4 * q) = o;
      }
      __syncthreads();
      int32_t* tmp = src; src = dst; dst = tmp;
    }
    const int4* s4 = reinterpret_cast<const int4*>(src);
    int4* g4 = reinterpret_cast<int4*>(gt);
    for (uint32_t i = threadIdx.x; i < tile_words / 4; i += THREADS) g4[i] = s4[i];
    __syncthreads();   // everyone done reading the buffers before the next bulk copy lands
  }
}

// x = n + n*(r + ir) of ONE generate sub-step, before ROUND (reference :99-102), for every owned cell;
// densities are not modified.  Diagnostic path for the "pre-rounding float update within 1e-6"
// criterion: one thread per (cell, species), operands straight from global memory, runtime tile width,
// the same operation order as the hot kernel.
__global__ void __launch_bounds__(256) pre_rounding_kernel(const GenParams p, uint32_t R, double* __restrict__ out) {
  const Layout& L = p.L;
  const unsigned long long n = (unsigned long long)L.nx * L.ny * L.nsp;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t s = (uint32_t)(i % L.nsp);
    const unsigned long long c = i / L.nsp;
    const uint32_t x = (uint32_t)(c % L.nx), yl = (uint32_t)(c / L.nx);
    const int32_t* tile = L.base + ((size_t)yl * L.ntx + x / L.tw) * L.tile_stride + x % L.tw;
    const int nd = tile[(size_t)s * L.tw];
    const double d = (double)(uint32_t)nd;
    const uint32_t b = s / R, rr = s % R;
    const uint32_t s0 = p.boff[b], W = p.boff[b + 1] - s0 - 1u;
    const uint4* blk = p.stream + (size_t)s0 * R + rr;
    const uint4 h = blk[0];
    const double diag = dbl(h.x, h.y), r = dbl(h.z, h.w);
    double ir = __dmul_rn(diag, d);
    for (uint32_t j = 0; j < W; ++j) {
      const uint4 w = blk[(size_t)(1 + j) * R];
      const double v = dbl(w.x, w.y);
      const int g = tile[w.z / 4];   // w.z = col*TW*4 bytes
      ir = __dadd_rn(ir, __dmul_rn(v, (double)(uint32_t)g));
    }
    out[i] = __dadd_rn(d, __dmul_rn(d, __dadd_rn(r, ir)));
  }
}

}  // namespace eco
