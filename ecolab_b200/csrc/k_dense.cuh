// k_dense.cuh -- generate() for a genuinely dense interaction matrix: the contraction  S[i,c] = sum_j beta[i,j] n[j,c]
// on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in TMEM), with the rest of
// EcolabPoint::generate (reference models/ecolab_model.cc:90-112) and ROUND (:57-66) fused into the epilogue.
//
// Exact-integer slicing (Ozaki-style): per species row i,  beta[i,j] * 2^s_i  is rounded to an integer F[i,j] of
// 8P-1 bits and written in balanced radix 256,  F = sum_p a_p 256^(P-1-p),  a_p in [-128,127]  (A operand, s8);
// a density is  n = sum_q b_q 256^q,  b_q in [0,255]  (B operand, u8).  Every slice product a_p x b_q is an
// s8 x u8 -> s32 GEMM whose K-sum is EXACT (8192 x 128 x 255 < 2^31 leaves room for 8 products per accumulator);
// products of equal weight 256^g, g = q-p+P-1, share one TMEM accumulator ("group").  The epilogue recombines
//     S = 2^-s_i * sum_g 256^g acc_g            (Horner in FP64: exact while the integer stays below 2^53, i.e. while
//                                                nsp * 2^(8P-1) * max n < 2^53; beyond that it rounds to 2^-53 relative)
//     ir = diag*n + S ;  x = n + n*(r + ir) ;  n' = ROUND(x)   with the same Philox draw as the sparse kernel.
// The only approximation is the rounding of beta to 8P-1 bits relative to its row maximum (P = 3: 2^-24, P = 7:
// beyond FP64), so the pre-rounding x agrees with the FP64 sequential sum to ~1e-9 (P = 3) and the integer results
// differ from the oracle only where that error flips a ROUND threshold; with beta entries that are dyadic rationals
// the path is bit-exact (tests/test_gpu_dense.py).  It cannot reproduce the reference's FP64 summation ORDER, which
// is why the sparse kernel stays the default and this path is selected only for dense beta (Engine::dense_wanted).
//
// Operands live in HBM as ready-made shared-memory images, so a pipeline stage is filled by two 1-D bulk copies
// (TMA unit, SASS UBLKCP) and needs no tensor map: K-major tiles of 128 rows x 128 bytes of K in the 128-byte swizzle
//     tile[row group = 16][row = 8][128 bytes],  16-byte chunk c of row r stored at chunk position c ^ (r % 8)
// for which the tcgen05 shared-memory descriptor is layout SWIZZLE_128B, SBO = 1024 (stride between 8-row groups), and a
// K step of 32 bytes advances the start address by 32 -- pinned on B200 by tools/umma_probe.cu
// (profiles/r1_umma_probe.txt).  Q = 2 density slices sit back to back and are multiplied as ONE N = 256 operand.
//     A: [m_tile][kc][p]  -> P consecutive tiles = one bulk copy        (built once per beta upload, host side)
//     B: [n_tile][kc][q]  -> Q consecutive tiles = one bulk copy        (dense_pack_kernel, every timestep)
// N runs over the padded cell quads of the slab in (row, x) order: N index = 4*(yl*nxq + xq) + e.
//
// Kernel = persistent, one CTA per SM, warp-specialised:
//   warp 0 : producer   -- bulk copies of stage (tile, kc) into a ring, "full"/"empty" mbarriers
//   warp 1 : MMA issuer -- one lane issues P*Q*(KC/32) tcgen05.mma per stage into <= 4 accumulators of 128 columns,
//                          tcgen05.commit frees the stage / publishes the accumulators
//   warps 2..17 : epilogue -- tcgen05.ld of the warp's 32 columns of every group (lane = species row), Horner into FP64 sums
//                            held in registers, accumulators handed back to the MMA warp at once; then the generate
//                            arithmetic, Philox ROUND and STG.128 run beside the next tile's MMAs (the densities of the next
//                            8 columns are prefetched meanwhile)
// Bound: shared-memory bandwidth -- per K chunk the bulk copies write (P + Q) x 16 KB and the MMAs read
// 4 KB of beta + N/128 x 4 KB of densities each, 160 KB per stage at P = 3, Q = 1 through 128 B/clk = 1250 cycles (768 tensor).
#pragma once
#include "k_common.cuh"
#include "k_generate.cuh"

namespace eco {

constexpr uint32_t DENSE_KC = 128;                   // K bytes (species) per pipeline stage and operand tile
constexpr uint32_t DENSE_TILE_BYTES = 128 * DENSE_KC;   // one 128-row operand tile
constexpr uint32_t DENSE_MAX_GROUPS = 4;             // 4 accumulators x 128 columns = the 512 TMEM columns
constexpr uint32_t DENSE_EPI_WARPS = 16;              // four per TMEM lane quadrant, 32 accumulator columns each
constexpr uint32_t DENSE_EPI_COLS = 128 / (DENSE_EPI_WARPS / 4);   // accumulator columns per epilogue warp
constexpr uint32_t DENSE_CHUNK = 8;                   // columns per tcgen05.ld (2 cell quads)
constexpr uint32_t DENSE_THREADS = (2 + DENSE_EPI_WARPS) * 32;

struct DenseParams {
  Layout L;                    // densities of the rows this launch covers
  const uint8_t* aop;          // beta slices (see above)
  const uint8_t* bop;          // density slices
  const double* diag;          // [mtiles*128] interaction.diag (0 in the padding)
  const double* rr;            // [mtiles*128] repro_rate
  const double* rscale;        // [mtiles*128] 2^-s_i
  double* partial;             // [ntn][mtiles][128 cols][128 rows] running sums between group rounds, or null
  unsigned int* err;           // watchdog word: set when a wait made no progress for ~10 s (checked by the host)
  unsigned long long* prof;    // optional [grid][8] cycle counters (tools only): MMA warp waits / issue, epilogue wait / work
  double* xout;                // cell-major [cell][nsp] pre-rounding x (diagnostic), or null
  uint32_t P, Q;               // slices of beta / of the densities
  uint32_t g0, g1;             // groups [g0, g1) handled by this launch (g1 - g0 <= 4)
  double w0;                   // 256^g0, the weight of the lowest group of this launch
  uint32_t flags;              // 1: add `partial` to this round's sum   2: store the sum to `partial`, no epilogue   4: per-stage cycle counters
  uint32_t nkc, mtiles, ntn;   // K chunks, 128-row species tiles, 128-column cell tiles
  uint32_t mblock;             // species tiles per L2 super-block (tile order: [m block][n tile][m in block])
  uint32_t nxq;                // padded quads per grid row = ntx*TW/4
  uint32_t nstages;
  unsigned long long t0;
  uint32_t rk[20];             // Philox key schedule (as GenParams)
};

// ---- tcgen05 wrappers ---------------------------------------------------------------------------------------------
// SWIZZLE_128B K-major descriptor: start address >> 4 in [0,14), LBO field 1 (unused for swizzled K-major), SBO = 1024 B
// in [32,46), version 1 in [46,48), layout type 2 in [61,64).  `lo` = (shared address >> 4) | 1 << 16.
constexpr uint32_t UMMA_DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint64_t umma_desc(uint32_t lo) { return ((uint64_t)UMMA_DESC_HI << 32) | lo; }
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
               "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
               : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {   // arrives on `bar` when all MMAs issued so far have completed
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// 16 consecutive 32-bit columns of this thread's TMEM lane (returns when the data is in the registers)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, int (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n\t"
      "tcgen05.wait::ld.sync.aligned;"   // same statement: no use of v[] can be scheduled before the wait
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, int (&v)[8]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
      : "r"(taddr)
      : "memory");
}

// mbarrier wait with a watchdog: a pipeline that stops making progress raises an error word instead of hanging the GPU (or
// trapping, which would kill the context under a debugger or time-slicing).  Once the word is set every wait of the launch
// returns at once, the kernel runs out, and the host finds the word at its next synchronisation and throws.
__device__ __forceinline__ void mbar_wait_wd(uint64_t* bar, uint32_t parity, unsigned int* err) {
  uint32_t done = 0;
  long long t0 = 0;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
    if (done) return;
    if (spin == 0) t0 = clock64();   // the clock is read only once a wait really waits (the MMA thread's common case does not)
    if ((spin & 1023u) == 1023u) {
      if (*reinterpret_cast<volatile unsigned int*>(err)) return;
      if (clock64() - t0 > 20000000000ll) { atomicExch(err, 1u); return; }   // ~10 s without progress
    }
  }
}

// tile id -> (species tile, cell tile); false for the padding of the last super-block
__device__ __forceinline__ bool dense_tile(const DenseParams& p, uint32_t id, uint32_t& mt, uint32_t& nt) {
  const uint32_t per = p.mblock * p.ntn;
  const uint32_t blk = id / per, rem = id - blk * per;
  nt = rem / p.mblock;
  mt = blk * p.mblock + (rem - nt * p.mblock);
  return mt < p.mtiles;
}

// ---- MMA issue of one pipeline stage, everything but the two base addresses known at compile time ------------------------
// A single thread issues the MMAs, so the tensor pipe runs only as fast as that thread gets through its instructions
// (measured: 64 cycles per M128 x N128 x 32 B MMA with a bare issue loop, ~300 with run-time slice loops): slice counts
// are template parameters, the accumulate flags and descriptor offsets fold to constants.  FIRST = first K step of a tile
// (accumulators start from zero; the N = 256 stacking of two density slices cannot be used there because a 256-column
// MMA has one accumulate flag for two groups that start at different moments).
constexpr uint32_t DENSE_IDESC128 = (2u << 4) | (1u << 7) | ((128u >> 3) << 17) | ((128u >> 4) << 24);   // D s32, A s8, B u8, K-major
constexpr uint32_t DENSE_IDESC256 = (2u << 4) | (1u << 7) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
template <int P, int Q, bool FIRST>
__device__ __forceinline__ void dense_issue_stage(uint32_t tmem, uint32_t a_lo, uint32_t b_lo) {
#pragma unroll
  for (int ks = 0; ks < (int)(DENSE_KC / 32); ++ks) {
#pragma unroll
    for (int pp = 0; pp < P; ++pp) {
      const uint64_t ad = umma_desc(a_lo + pp * (DENSE_TILE_BYTES >> 4) + ks * 2);
      if (Q == 2 && !(FIRST && ks == 0)) {
        umma_i8(tmem + (P - 1 - pp) * 128, ad, umma_desc(b_lo + ks * 2), DENSE_IDESC256, 1u);
      } else {
#pragma unroll
        for (int q = 0; q < Q; ++q)   // group g = q + P-1-pp was touched before by (pp-1, q-1) when both exist
          umma_i8(tmem + (q + P - 1 - pp) * 128, ad, umma_desc(b_lo + q * (DENSE_TILE_BYTES >> 4) + ks * 2), DENSE_IDESC128,
                  (FIRST && ks == 0 && !(pp > 0 && q > 0)) ? 0u : 1u);
      }
    }
  }
}

// The same for ONE ROUND of a slicing with more than four weight groups (P + Q - 1 > 4: densities of 2^16 and more with three
// beta slices): only the slice pairs whose group lies in [G0, G1) are issued, into accumulators numbered from G0.  Everything
// is still a compile-time constant, so the issuing thread runs at the tensor pipe's pace instead of the ~300 cycles per MMA
// of run-time slice loops (a Q = 3 step took 87 ms through those).
template <int P, int Q, int G0, int G1, bool FIRST>
__device__ __forceinline__ void dense_issue_stage_range(uint32_t tmem, uint32_t a_lo, uint32_t b_lo) {
#pragma unroll
  for (int ks = 0; ks < (int)(DENSE_KC / 32); ++ks) {
#pragma unroll
    for (int pp = 0; pp < P; ++pp) {
      const uint64_t ad = umma_desc(a_lo + pp * (DENSE_TILE_BYTES >> 4) + ks * 2);
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int g = q + P - 1 - pp;
        if (g < G0 || g >= G1) continue;
        // group g was touched before, in this K step, by (pp-1, q-1) when both exist (same g, earlier in the loop order)
        umma_i8(tmem + (g - G0) * 128, ad, umma_desc(b_lo + q * (DENSE_TILE_BYTES >> 4) + ks * 2), DENSE_IDESC128,
                (FIRST && ks == 0 && !(pp > 0 && q > 0)) ? 0u : 1u);
      }
    }
  }
}

// ---- B operand: int32 densities -> Q u8 slices in the shared-memory image layout ---------------------------------------
// One thread = 16 consecutive species x one cell quad: 16 LDG.128 (contiguous 256 B when TW = 4) -> 4*Q STG.128.
__global__ void __launch_bounds__(256) dense_pack_kernel(const Layout L, uint8_t* __restrict__ bop, uint32_t Q, uint32_t nkc,
                                                         uint32_t ntn, uint32_t nxq, unsigned int* __restrict__ maxn) {
  const uint32_t k16_total = nkc * (DENSE_KC / 16);
  const unsigned long long nquads = (unsigned long long)ntn * 32;   // padded: 32 quads per cell tile
  const unsigned long long nwork = nquads * k16_total;
  const uint32_t qpt = L.tw / 4;
  uint32_t mx = 0;
  for (unsigned long long w = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; w < nwork;
       w += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t kg = (uint32_t)(w % k16_total);
    const unsigned long long qi = w / k16_total;
    const uint32_t yl = (uint32_t)(qi / nxq), xq = (uint32_t)(qi - (unsigned long long)yl * nxq);
    uint32_t v[16][4];
    const uint32_t j0 = kg * 16;
    if (yl < L.ny) {
      const int32_t* t = L.base + ((size_t)yl * L.ntx + xq / qpt) * L.tile_stride + 4 * (xq % qpt);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        int4 d = make_int4(0, 0, 0, 0);
        if (j0 + i < L.cap) d = __ldg(reinterpret_cast<const int4*>(t + (size_t)(j0 + i) * L.tw));
        v[i][0] = (uint32_t)d.x; v[i][1] = (uint32_t)d.y; v[i][2] = (uint32_t)d.z; v[i][3] = (uint32_t)d.w;
        mx = max(mx, max(max(v[i][0], v[i][1]), max(v[i][2], v[i][3])));
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i][0] = v[i][1] = v[i][2] = v[i][3] = 0;
    }
    const uint32_t nt = (uint32_t)(qi / 32), ql = (uint32_t)(qi % 32);   // cell tile, quad inside it
    const uint32_t kc = kg / (DENSE_KC / 16), k16 = kg % (DENSE_KC / 16);
    for (uint32_t q = 0; q < Q; ++q) {
      // rows 4*ql .. 4*ql+3 of the tile; row r = 8*(r/8) + r%8 lives at (r/8)*1024 + (r%8)*128, its chunk k16 at (k16 ^ r%8)*16
      uint8_t* dst = bop + (((size_t)nt * nkc + kc) * Q + q) * DENSE_TILE_BYTES + (size_t)(ql >> 1) * 1024;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const uint32_t r8 = (ql & 1u) * 4 + e;
        uint32_t wds[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          wds[k] = ((v[4 * k][e] >> (8 * q)) & 255u) | (((v[4 * k + 1][e] >> (8 * q)) & 255u) << 8) |
                   (((v[4 * k + 2][e] >> (8 * q)) & 255u) << 16) | (((v[4 * k + 3][e] >> (8 * q)) & 255u) << 24);
        *reinterpret_cast<uint4*>(dst + r8 * 128 + ((k16 ^ r8) * 16)) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
      }
    }
  }
  mx = __reduce_max_sync(FULL, mx);
  if ((threadIdx.x & 31u) == 0 && mx) atomicMax(maxn, mx);
}

// ---- the fused dense generate kernel -----------------------------------------------------------------------------------
// NGR = weight groups of this launch when it is a plain single-round launch (flags == 0): the recombination unrolls
// without predicates; NGR = 0 is the general form (any group count, partial sums between rounds).
template <int NGR>
__global__ void __launch_bounds__(DENSE_THREADS, 1) generate_dense_kernel(const DenseParams p) {
  extern __shared__ __align__(1024) unsigned char dense_smem[];
  __shared__ uint64_t s_full[8], s_empty[8], s_tfull, s_tempty;
  __shared__ uint32_t s_tmem;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t S = p.nstages, P = p.P, Q = p.Q;
  const uint32_t a_bytes = P * DENSE_TILE_BYTES, b_bytes = Q * DENSE_TILE_BYTES, stage_bytes = a_bytes + b_bytes;
  const uint32_t ntile_ids = ((p.mtiles + p.mblock - 1) / p.mblock) * p.mblock * p.ntn;

  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < S; ++s) { mbar_init(s_full + s, 1); mbar_init(s_empty + s, 1); }
    mbar_init(&s_tfull, 1);
    mbar_init(&s_tempty, DENSE_EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  if (warp == 0) {
    // ===================== producer =====================
    if (lane == 0) {
      uint32_t stage = 0, round = 0;
      for (uint32_t id = blockIdx.x; id < ntile_ids; id += gridDim.x) {
        uint32_t mt, nt;
        if (!dense_tile(p, id, mt, nt)) continue;
        const uint8_t* a = p.aop + (size_t)mt * p.nkc * a_bytes;
        const uint8_t* b = p.bop + (size_t)nt * p.nkc * b_bytes;
        for (uint32_t kc = 0; kc < p.nkc; ++kc) {
          if (round) mbar_wait_wd(s_empty + stage, (round - 1u) & 1u, p.err);
          unsigned char* dst = dense_smem + (size_t)stage * stage_bytes;
          mbar_expect_tx(s_full + stage, stage_bytes);
          bulk_g2s(dst, a + (size_t)kc * a_bytes, a_bytes, s_full + stage);
          bulk_g2s(dst + a_bytes, b + (size_t)kc * b_bytes, b_bytes, s_full + stage);
          if (++stage == S) { stage = 0; ++round; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      long long pc[3] = {0, 0, 0};
      // per-stage cycle counters cost the ONE issuing thread three clock reads per stage: only on request (flags & 4, tools);
      // otherwise prof[2] is the cycle count of the whole loop, which with prof[7] (its wall time) gives the kernel's real SM clock
      const bool prof = p.prof != nullptr && (p.flags & 4u);
      const long long cstart = p.prof ? clock64() : 0;
      unsigned long long ns0 = 0;
      if (p.prof) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns0));
      // compile-time issue paths: single-round launches of the common slicings, and the two rounds of P = 3 with Q = 3 / 4
      const uint32_t fast = (p.g0 == 0 && p.g1 == P + Q - 1u) ? P * 10u + Q
                            : (P == 3u && (Q == 3u || Q == 4u) && (p.g0 == 0u || p.g0 == 4u)) ? 1000u + Q * 10u + p.g0 : 0u;
      uint32_t stage = 0, round = 0, k = 0;
      for (uint32_t id = blockIdx.x; id < ntile_ids; id += gridDim.x) {
        uint32_t mt, nt;
        if (!dense_tile(p, id, mt, nt)) continue;
        long long c0 = prof ? clock64() : 0;
        if (k) mbar_wait_wd(&s_tempty, (k - 1u) & 1u, p.err);   // the epilogue has drained the accumulators of the previous tile
        tc_fence_after();
        if (prof) pc[0] += clock64() - c0;
        uint32_t written = 0;                             // groups that already hold a partial sum of this tile
        for (uint32_t kc = 0; kc < p.nkc; ++kc) {
          if (prof) c0 = clock64();
          mbar_wait_wd(s_full + stage, round & 1u, p.err);
          tc_fence_after();
          const long long c1 = prof ? clock64() : 0;
          if (prof) pc[1] += c1 - c0;
          const uint32_t sa = smem_u32(dense_smem + (size_t)stage * stage_bytes);
          const uint32_t a_lo = (sa >> 4) | (1u << 16), b_lo = ((sa + a_bytes) >> 4) | (1u << 16);
          if (fast == 31u) { if (kc) dense_issue_stage<3, 1, false>(tmem, a_lo, b_lo); else dense_issue_stage<3, 1, true>(tmem, a_lo, b_lo); }
          else if (fast == 32u) { if (kc) dense_issue_stage<3, 2, false>(tmem, a_lo, b_lo); else dense_issue_stage<3, 2, true>(tmem, a_lo, b_lo); }
          else if (fast == 21u) { if (kc) dense_issue_stage<2, 1, false>(tmem, a_lo, b_lo); else dense_issue_stage<2, 1, true>(tmem, a_lo, b_lo); }
          else if (fast == 22u) { if (kc) dense_issue_stage<2, 2, false>(tmem, a_lo, b_lo); else dense_issue_stage<2, 2, true>(tmem, a_lo, b_lo); }
          else if (fast == 41u) { if (kc) dense_issue_stage<4, 1, false>(tmem, a_lo, b_lo); else dense_issue_stage<4, 1, true>(tmem, a_lo, b_lo); }
          else if (fast == 1030u) { if (kc) dense_issue_stage_range<3, 3, 0, 4, false>(tmem, a_lo, b_lo); else dense_issue_stage_range<3, 3, 0, 4, true>(tmem, a_lo, b_lo); }
          else if (fast == 1034u) { if (kc) dense_issue_stage_range<3, 3, 4, 5, false>(tmem, a_lo, b_lo); else dense_issue_stage_range<3, 3, 4, 5, true>(tmem, a_lo, b_lo); }
          else if (fast == 1040u) { if (kc) dense_issue_stage_range<3, 4, 0, 4, false>(tmem, a_lo, b_lo); else dense_issue_stage_range<3, 4, 0, 4, true>(tmem, a_lo, b_lo); }
          else if (fast == 1044u) { if (kc) dense_issue_stage_range<3, 4, 4, 6, false>(tmem, a_lo, b_lo); else dense_issue_stage_range<3, 4, 4, 6, true>(tmem, a_lo, b_lo); }
          else {
            // any other slicing / a multi-round launch: run-time loops (correct, slower to issue)
#pragma unroll
            for (uint32_t ks = 0; ks < DENSE_KC / 32; ++ks)
              for (uint32_t pp = 0; pp < P; ++pp)
                for (uint32_t q = 0; q < Q; ++q) {
                  const uint32_t g = q + (P - 1u) - pp;
                  if (g < p.g0 || g >= p.g1) continue;
                  umma_i8(tmem + (g - p.g0) * 128u, umma_desc(a_lo + pp * (DENSE_TILE_BYTES >> 4) + ks * 2u),
                          umma_desc(b_lo + q * (DENSE_TILE_BYTES >> 4) + ks * 2u), DENSE_IDESC128, (written >> g) & 1u);
                  written |= 1u << g;
                }
          }
          umma_commit(s_empty + stage);                   // stage reusable once these MMAs have read it
          if (prof) pc[2] += clock64() - c1;
          if (++stage == S) { stage = 0; ++round; }
        }
        umma_commit(&s_tfull);                            // accumulators complete
        ++k;
      }
      if (p.prof && !prof) pc[2] = clock64() - cstart;
      if (p.prof) { p.prof[blockIdx.x * 8 + 0] = pc[0]; p.prof[blockIdx.x * 8 + 1] = pc[1]; p.prof[blockIdx.x * 8 + 2] = pc[2]; p.prof[blockIdx.x * 8 + 3] = k;
        unsigned long long ns1;   // wall time of the issue loop: cycles / ns = the SM clock the kernel really ran at
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
        p.prof[blockIdx.x * 8 + 7] = ns1 - ns0; }
    }
  } else {
    // ===================== epilogue (warps 2..17; TMEM lanes 32*(warp%4).., columns DENSE_EPI_COLS*((warp-2)/4)..) =====================
    const uint32_t lq = warp & 3u, row = lq * 32u + lane, cbase = ((warp - 2u) >> 2) * DENSE_EPI_COLS;
    const uint32_t ngr = NGR ? (uint32_t)NGR : p.g1 - p.g0;
    const uint32_t qsh = p.L.tw == 4 ? 0u : p.L.tw == 8 ? 1u : p.L.tw == 16 ? 2u : 3u;   // log2(quads per density tile)
    const double w0 = p.w0;
    const bool final_round = NGR ? true : !(p.flags & 2u), add_partial = NGR ? false : (p.flags & 1u) != 0;
    long long ew = 0, ed = 0, ex = 0;   // cycles: waiting for the accumulators, draining them, the update arithmetic
    uint32_t k = 0;
    // a cursor walks the padded quads of the slab in (row, x) order; null pointer = nothing to update there
    struct Cursor { uint32_t yl, xq; };
    auto advance = [&](Cursor& c) { if (++c.xq == p.nxq) { c.xq = 0; ++c.yl; } };
    for (uint32_t id = blockIdx.x; id < ntile_ids; id += gridDim.x) {
      uint32_t mt, nt;
      if (!dense_tile(p, id, mt, nt)) continue;
      const uint32_t sp = mt * 128u + row;
      const double diag = p.diag[sp], rr = p.rr[sp], rs = p.rscale[sp];
      double* part = p.partial ? p.partial + ((size_t)nt * p.mtiles + mt) * (128 * 128) + row : nullptr;
      const bool live = final_round && sp < p.L.nsp;
      int32_t* const sp_base = p.L.base + (size_t)sp * p.L.tw;
      auto quad_ptr = [&](const Cursor& c) -> int32_t* {
        if (!live || c.yl >= p.L.ny) return nullptr;
        return sp_base + ((size_t)c.yl * p.L.ntx + (c.xq >> qsh)) * p.L.tile_stride + 4u * (c.xq & ((1u << qsh) - 1u));
      };
      const uint32_t q0 = nt * 32u + (cbase >> 2);   // first quad of this warp's columns
      Cursor cw, cp;
      cw.yl = q0 / p.nxq; cw.xq = q0 - cw.yl * p.nxq;
      cp = cw;
      constexpr uint32_t NQ = DENSE_CHUNK / 4;   // quads per chunk
      int4 own[NQ], nxt[NQ];
#pragma unroll
      for (uint32_t qd = 0; qd < NQ; ++qd) {
        const int32_t* gp = quad_ptr(cp);
        nxt[qd] = gp ? *reinterpret_cast<const int4*>(gp) : make_int4(0, 0, 0, 0);
        advance(cp);
      }
      const long long e0 = clock64();
      mbar_wait_wd(&s_tfull, k & 1u, p.err);
      tc_fence_after();
      const long long e1 = clock64();
      ew += e1 - e0;
      // Drain first: all of this warp's columns leave TMEM as recombined FP64 sums held in registers, the accumulators go
      // back to the MMA warp at once, and the update arithmetic below overlaps the next tile's MMAs.
      constexpr uint32_t NCH = DENSE_EPI_COLS / DENSE_CHUNK;
      double hsum[NCH][DENSE_CHUNK];
#pragma unroll
      for (uint32_t ci = 0; ci < NCH; ++ci) {
        int acc[DENSE_MAX_GROUPS][DENSE_CHUNK];
#pragma unroll
        for (uint32_t g = 0; g < DENSE_MAX_GROUPS; ++g)
          if (g < ngr) tmem_ld8(tmem + ((lq * 32u) << 16) + g * 128u + cbase + ci * DENSE_CHUNK, acc[g]);
#pragma unroll
        for (uint32_t j = 0; j < DENSE_CHUNK; ++j) {
          // Horner over the weight groups.  fma is used ONLY here; on integers below 2^53 t*256 + acc is exact, so the
          // fused and the two-step form give the same bits (the reference-order arithmetic below stays unfused).
          double t = 0.0;
          bool started = false;
#pragma unroll
          for (int g = DENSE_MAX_GROUPS - 1; g >= 0; --g)
            if ((uint32_t)g < ngr) {
              t = started ? __fma_rn(t, 256.0, (double)acc[g][j]) : (double)acc[g][j];
              started = true;
            }
          hsum[ci][j] = t;
        }
      }
#ifndef ECO_DENSE_LATE_RELEASE
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_tempty);
#endif
      const long long e2 = clock64();
      ed += e2 - e1;
#pragma unroll
      for (uint32_t ci = 0; ci < NCH; ++ci) {
        const uint32_t c0 = cbase + ci * DENSE_CHUNK;
#pragma unroll
        for (uint32_t qd = 0; qd < NQ; ++qd) own[qd] = nxt[qd];
        if (ci + 1 < NCH) {
#pragma unroll
          for (uint32_t qd = 0; qd < NQ; ++qd) {
            const int32_t* gp = quad_ptr(cp);
            nxt[qd] = gp ? *reinterpret_cast<const int4*>(gp) : make_int4(0, 0, 0, 0);
            advance(cp);
          }
        }
#pragma unroll
        for (uint32_t qd = 0; qd < NQ; ++qd) {
          double sum[4];
#pragma unroll
          for (uint32_t e = 0; e < 4; ++e) {
            double t = hsum[ci][4 * qd + e];
            if (!NGR && w0 != 1.0) t = __dmul_rn(t, w0);
            if (add_partial) t = __dadd_rn(t, part[(size_t)(c0 + 4 * qd + e) * 128]);
            sum[e] = t;
          }
          const Cursor c = cw;
          advance(cw);
          if (!final_round) {
#pragma unroll
            for (uint32_t e = 0; e < 4; ++e) part[(size_t)(c0 + 4 * qd + e) * 128] = sum[e];
            continue;
          }
          int32_t* gp = quad_ptr(c);
          if (!gp) continue;
          const int4 o4 = own[qd];
          const double d[4] = {(double)(uint32_t)o4.x, (double)(uint32_t)o4.y, (double)(uint32_t)o4.z, (double)(uint32_t)o4.w};
          double x[4];
#pragma unroll
          for (uint32_t e = 0; e < 4; ++e) {
            const double ir = __dadd_rn(__dmul_rn(diag, d[e]), __dmul_rn(sum[e], rs));
            x[e] = __dadd_rn(d[e], __dmul_rn(d[e], __dadd_rn(rr, ir)));
          }
          if (p.xout) {
#pragma unroll
            for (uint32_t e = 0; e < 4; ++e) {
              const uint32_t xc = c.xq * 4 + e;
              if (xc < p.L.nx) p.xout[((size_t)c.yl * p.L.nx + xc) * p.L.nsp + sp] = x[e];
            }
            continue;
          }
          const uint4 u = philox_rk(c.xq, p.L.y0 + c.yl, sp, (uint32_t)p.t0, p.rk);
          // ROUND: in-range fast path decided per thread on the high words (negative x has the sign bit set -> huge)
          const uint32_t hm = max(max((uint32_t)__double2hiint(x[0]), (uint32_t)__double2hiint(x[1])),
                                  max((uint32_t)__double2hiint(x[2]), (uint32_t)__double2hiint(x[3])));
          int4 o;
          if (hm >= 0x41DFFFFFu) {
            o.x = round_stoch(x[0], u.x); o.y = round_stoch(x[1], u.y);
            o.z = round_stoch(x[2], u.z); o.w = round_stoch(x[3], u.w);
          } else {
            o.x = round_stoch_inrange(x[0], u.x); o.y = round_stoch_inrange(x[1], u.y);
            o.z = round_stoch_inrange(x[2], u.z); o.w = round_stoch_inrange(x[3], u.w);
          }
          *reinterpret_cast<int4*>(gp) = o;
        }
      }
#ifdef ECO_DENSE_LATE_RELEASE   // A/B build (tools/variants.sh): hold the accumulators until the update is stored, as round 1 did
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_tempty);
#endif
      ex += clock64() - e2;
      ++k;
    }
    if (p.prof && warp == 2 && lane == 0) { p.prof[blockIdx.x * 8 + 4] = ew; p.prof[blockIdx.x * 8 + 5] = ex; p.prof[blockIdx.x * 8 + 6] = ed; }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u));
  }
}

}  // namespace eco
