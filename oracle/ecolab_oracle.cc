// ecolab_oracle.cc -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the EcoLab model's per-timestep population update
// (generate / mutate / migrate / condense) used as the parity oracle for the
// CUDA path and as the timed CPU baseline in bench.py.  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load this library.  The product (ecolab_b200/) never links or calls it.
//
// The reference (highperformancecoder/ecolab) cannot be compiled here: every
// translation unit needs the absent `classdesc` submodule.  The arithmetic of
// the hot path, however, lives entirely in files that ARE present, and this
// file restates it function by function (reference file:line cited at each
// function; paths relative to the reference checkout):
//     models/ecolab_model.cc, models/ecolab_model.h,
//     include/arrays.h, include/sparse_mat.h, src/arrays.cc, src/random.cc
//
// Parity pin: in RNG mode MT19937 this restatement reproduces the reference's
// golden vector test/00/predator-prey.sh:56-67 bit-for-bit (tests/
// test_oracle_golden.py), plus test/00/condense.sh, the conservation property
// of test/00/spatial_ecolab.sh and the primitives pinned by
// test/testarrays.cc:284-327.  ModelData::mutate's numerics are pinned by no
// reference test ("parity unpinned" for that one function; statistical +
// same-libc-stream comparison only).
//
// Two RNG back ends behind one draw():
//   MT19937 : one std::mt19937 per cell, consumed sequentially exactly as the
//             reference does (models/ecolab_model.h:97-102).
//   PHILOX  : counter-based Philox4x32-10, the stream shared with the GPU
//             (contract in DESIGN.md section "RNG contract"), cell (x, y) = (id % numX, id / numX):
//               ctr = {(x>>2) | stage<<30, y, slot, t_lo}, key = {seed, t_hi}
//               draw = out[x & 3]
//
// Build: g++ -O3 -fopenmp -ffp-contract=off (the reference builds x86-64
// without FMA, include/Makefile:567-571); see oracle/Makefile.

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

typedef double Float;  // models/ecolab_model.h:18-22 (USE_FLOAT off)

enum RngMode { RNG_MT19937 = 0, RNG_PHILOX = 1 };
enum Stage { ST_GENERATE = 0, ST_MUTATE = 1, ST_SALT = 2 };

// ---- Philox4x32-10 (Salmon et al., SC'11), restated independently of csrc/ ----
inline void philox_round(uint32_t c[4], const uint32_t k[2]) {
  const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
  const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
  const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k[0];
  const uint32_t n1 = (uint32_t)p1;
  const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k[1];
  const uint32_t n3 = (uint32_t)p0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
inline void philox_block(uint32_t c[4], uint32_t k0, uint32_t k1) {
  uint32_t k[2] = {k0, k1};
  for (int r = 0; r < 10; ++r) {
    if (r) { k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u; }
    philox_round(c, k);
  }
}
// the shared stream: one block serves the 4 consecutive cells x&~3 .. x|3 of grid row y for one species slot
inline uint32_t philox_draw(uint32_t seed, uint32_t stage, uint32_t x, uint32_t y,
                            uint32_t slot, uint64_t t) {
  uint32_t c[4] = {(x >> 2) | (stage << 30), y, slot, (uint32_t)t};
  philox_block(c, seed, (uint32_t)(t >> 32));
  return c[x & 3];
}

// ---- libc-rand() based generators used by ModelData::mutate --------------------
// src/random.cc:57 (urand::rand, neither UNURAN nor GNUSL)
inline double urand() { return ((double)::rand()) / RAND_MAX; }
// src/random.cc:59-64 (gaussrand::rand: a *running* CLT generator); the
// reference's array_grand is one process-wide object (src/arrays.cc:25).
struct GaussRand {
  double sum = 0; unsigned n = 0;
  double rand() { sum += urand(); n++; return (sum - .5 * n) * std::sqrt(12. / n); }
} g_array_grand;

struct Cell {
  uint32_t id = 0;
  std::vector<int32_t> n;       // EcolabPoint::density
  std::mt19937 mt;              // EcolabPoint::rand
  Float salt = 0;
  std::vector<uint32_t> nbr;    // indices into Oracle::cells
};

struct Oracle {
  // ModelData, models/ecolab_model.h:39-68
  std::vector<int32_t> species;
  std::vector<Float> create, repro_rate, mutation, migration;
  std::vector<Float> diag, val;
  std::vector<uint32_t> row, col;
  std::vector<std::vector<uint32_t>> rowlist;   // oDiagIdx
  uint64_t tstep = 0, last_mut_tstep = 0, last_mig_tstep = 0;
  Float sp_sep = 0.1, mut_max = 0, repro_min = 0, repro_max = 1, odiag_min = 0,
        odiag_max = 1, gen_bias = 0;
  bool fixMutation = false, fixMigration = false;
  // SpatialModel, models/ecolab_model.h:128-151
  size_t numX = 1, numY = 1, maxNbrs = 0;
  bool panmictic = true;
  std::vector<Cell> cells;
  // RNG
  int rng_mode = RNG_MT19937;
  uint32_t seed = 0;
  int sp_cntr = -1;   // function-static in the reference (models/ecolab_model.cc:366)
  unsigned nprocs = 1;
  // scratch for the float-tolerance test: x before ROUND of the last generate sub-step
  bool record_pre = false;
  std::vector<Float> pre;   // [cell][nsp]
  // row-slab view used by the multi-rank CPU tests: this instance holds global rows [y0, y0+nyl)
  // of the numX x numY grid; ghost rows carry the neighbouring slabs' boundary rows for migrate
  size_t y0 = 0, nyl = 1;
  std::vector<int32_t> ghost_top, ghost_bottom;   // [numX][nsp]
  struct Rec { uint32_t cell, slot, count; };
  std::vector<Rec> pending;

  Oracle() { cells.resize(1); }   // panmictic: the model *is* one cell (models/ecolab_model.h:105)

  size_t nsp() const { return species.size(); }

  uint32_t draw(Cell& c, uint32_t stage, uint32_t slot, uint64_t t) {
    if (rng_mode == RNG_MT19937) return (uint32_t)c.mt();
    return philox_draw(seed, stage, (uint32_t)(c.id % numX), (uint32_t)(c.id / numX), slot, t);
  }

  // models/ecolab_model.cc:57-66  EcolabPoint::ROUND
  int ROUND(Float x, Cell& c, uint32_t stage, uint32_t slot, uint64_t t) {
    const Float maxInt = Float(INT_MAX - 1);
    if (x <= 0) return 0;
    if (x >= maxInt) return (int)maxInt;
    Float ip;
    const Float frac = std::fabs(std::modf(x, &ip));
    const uint32_t u = draw(c, stage, slot, t);   // exactly one draw, middle branch only
    return frac * Float(4294967295u) > Float(u) ? (int)x + 1 : (int)x;
  }

  // models/ecolab_model.cc:718-723  ModelData::computeODiagIdx
  void computeODiagIdx() {
    rowlist.assign(nsp(), std::vector<uint32_t>());
    for (uint32_t j = 0; j < row.size(); ++j) rowlist[row[j]].push_back(j);
  }

  // models/ecolab_model.cc:705-716 ModelData::makeConsistent and :725-735 SpatialModel::makeConsistent
  void makeConsistent() {
    size_t n = nsp();
    for (auto& c : cells) n = std::max(n, c.n.size());
    if (species.empty()) { species.resize(n); for (size_t i = 0; i < n; ++i) species[i] = (int)i; }
    if (create.empty()) create.assign(nsp(), 0);
    if (mutation.empty()) mutation.assign(nsp(), 0);
    if (migration.empty()) migration.assign(nsp(), 0);
    computeODiagIdx();
  }

  // models/ecolab_model.cc:580-605 SpatialModel::setGrid (open boundaries, id = x + nx*y)
  void setGrid(size_t nx, size_t ny) {
    panmictic = false;
    numX = nx; numY = ny;
    cells.assign(nx * ny, Cell());
    for (size_t i = 0; i < nx; ++i)
      for (size_t j = 0; j < ny; ++j) {
        Cell& c = cells[i + nx * j];
        c.id = (uint32_t)(i + nx * j);
        c.mt.seed(c.id);
        if (i > 0) c.nbr.push_back((uint32_t)(i - 1 + nx * j));
        if (i < nx - 1) c.nbr.push_back((uint32_t)(i + 1 + nx * j));
        if (j > 0) c.nbr.push_back((uint32_t)(i + nx * (j - 1)));
        if (j < ny - 1) c.nbr.push_back((uint32_t)(i + nx * (j + 1)));
      }
    for (auto& c : cells) maxNbrs = std::max(maxNbrs, c.nbr.size());   // never shrinks
  }

  // The same grid, but only rows [y0_, y0_+nyl_) live here (cell ids stay global: x + nx*y).
  // In-slab neighbours as in setGrid; the rows above/below come from ghost_top/ghost_bottom.
  void setGridSlab(size_t nx, size_t ny, size_t y0_, size_t nyl_) {
    panmictic = false;
    numX = nx; numY = ny; y0 = y0_; nyl = nyl_;
    cells.assign(nx * nyl, Cell());
    for (size_t j = 0; j < nyl; ++j)
      for (size_t i = 0; i < nx; ++i) {
        Cell& c = cells[i + nx * j];
        c.id = (uint32_t)(i + nx * (y0 + j));
        c.mt.seed(c.id);
      }
    const size_t nb = std::min<size_t>(nx - 1, 2) + std::min<size_t>(ny - 1, 2);
    maxNbrs = std::max(maxNbrs, nb);
  }

  // per-rank half of migrate (:631-664) on a slab: neighbours W,E,N,S in the reference's order
  // (:593-600: x-1, x+1, y-1, y+1), rows outside the slab read from the ghost rows
  uint64_t migrateLocal() {
    const size_t ns = nsp();
    std::vector<Float> cm(ns);
    const Float cap = 1.0 / maxNbrs;
    for (size_t s = 0; s < ns; ++s) {
      const Float mm = (tstep - last_mig_tstep) * migration[s];
      cm[s] = mm > cap ? cap : mm;
    }
    std::vector<std::vector<int32_t>> delta(cells.size(), std::vector<int32_t>(ns, 0));
    auto nbr = [&](long x, long yl) -> const int32_t* {   // yl relative to the slab
      if (yl < 0) return ghost_top.data() + (size_t)x * ns;
      if (yl >= (long)nyl) return ghost_bottom.data() + (size_t)x * ns;
      return cells[x + numX * yl].n.data();
    };
    for (long yl = 0; yl < (long)nyl; ++yl)
      for (long x = 0; x < (long)numX; ++x) {
        const size_t ci = x + numX * yl;
        const long yg = (long)y0 + yl;
        const int32_t* own = cells[ci].n.data();
        const int32_t* nb[4] = {x > 0 ? nbr(x - 1, yl) : nullptr, x < (long)numX - 1 ? nbr(x + 1, yl) : nullptr,
                                yg > 0 ? nbr(x, yl - 1) : nullptr, yg < (long)numY - 1 ? nbr(x, yl + 1) : nullptr};
        for (int k = 0; k < 4; ++k)
          if (nb[k])
            for (size_t s = 0; s < ns; ++s) delta[ci][s] += (int32_t)(cm[s] * (nb[k][s] - own[s]));
      }
    uint64_t total = 0;
    for (size_t ci = 0; ci < cells.size(); ++ci)
      for (size_t s = 0; s < ns; ++s) { cells[ci].n[s] += delta[ci][s]; total += (uint64_t)std::abs(delta[ci][s]); }
    last_mig_tstep = tstep;
    return total;
  }

  // per-rank halves of condense (:173-179 mask, :188-194 apply)
  void aliveMask(uint8_t* mask) {
    const size_t ns = nsp();
    for (size_t s = 0; s < ns; ++s) mask[s] = 0;
    for (auto& c : cells)
      for (size_t s = 0; s < c.n.size() && s < ns; ++s) if (c.n[s]) mask[s] = 1;
  }
  unsigned condenseApply(const uint8_t* m) {
    const size_t ns = nsp();
    std::vector<char> mask(m, m + ns);
    size_t ntrue = 0; for (char b : mask) ntrue += b ? 1 : 0;
    if (ntrue == ns) return 0;
    condenseModel(mask);
    for (auto& c : cells) {
      std::vector<int32_t> r;
      for (size_t s = 0; s < c.n.size(); ++s) if (mask[s] && r.size() < ntrue) r.push_back(c.n[s]);
      c.n.swap(r);
    }
    return (unsigned)(ns - ntrue);
  }

  // per-rank halves of mutate: :224-227 (cells), then the density append :272-275
  void mutateCells() {
    const size_t ns = nsp();
    std::vector<Float> ms(ns);
    for (size_t s = 0; s < ns; ++s)
      ms[s] = sp_sep * repro_rate[s] * mutation[s] * Float(tstep - last_mut_tstep);
    last_mut_tstep = tstep;
    pending.clear();
    for (auto& c : cells) {
      if (c.n.empty()) continue;
      std::vector<uint32_t> spec(c.n.size());
      uint64_t tot = 0;
      for (size_t s = 0; s < c.n.size(); ++s) { spec[s] = (uint32_t)ROUND(ms[s] * c.n[s], c, ST_MUTATE, (uint32_t)s, tstep); tot += spec[s]; }
      if (!tot) continue;
      for (size_t s = 0; s < c.n.size(); ++s)
        if (spec[s]) { c.n[s] -= (int32_t)spec[s]; pending.push_back(Rec{c.id, (uint32_t)s, spec[s]}); }
    }
  }
  void appendSpecies(const uint32_t* ids, size_t m) {
    for (auto& c : cells)
      for (size_t k = 0; k < m; ++k) c.n.push_back(ids[k] == c.id ? 1 : 0);
  }

  // models/ecolab_model.cc:90-112 EcolabPoint::generate; :607-614 / :49-54 callers
  void generate(unsigned niter) {
    if (tstep == 0) makeConsistent();
    const size_t ns = nsp();
    if (rowlist.size() != ns) computeODiagIdx();   // a model loaded mid-run (tstep set by hand) never went through makeConsistent
    if (record_pre) pre.assign(cells.size() * ns, 0);
    const long ncell = (long)cells.size();
#pragma omp parallel for schedule(static)
    for (long ci = 0; ci < ncell; ++ci) {
      Cell& c = cells[ci];
      std::vector<int32_t> l(c.n), t(c.n.size());
      for (unsigned step = 0; step < niter; ++step) {
        for (size_t i = 0; i < l.size(); ++i) {
          Float ir = diag[i] * l[i];
          for (uint32_t j : rowlist[i]) ir += val[j] * l[col[j]];
          const Float x = l[i] + l[i] * (repro_rate[i] + ir);
          if (record_pre && step + 1 == niter) pre[ci * ns + i] = x;
          t[i] = ROUND(x, c, ST_GENERATE, (uint32_t)i, tstep + step);
        }
        l.swap(t);
      }
      c.n = l;
    }
    tstep += niter;
  }

  // models/ecolab_model.cc:616-703 SpatialModel::migrate
  uint64_t migrate() {
    const long ncell = (long)cells.size();
    for (auto& c : cells)   // :619-620, the (unused) salt draw
      c.salt = Float(draw(c, ST_SALT, 0, tstep)) / Float(4294967295u);
    const size_t ns = nsp();
    std::vector<Float> cm(ns);
    const Float cap = 1.0 / maxNbrs;
    for (size_t s = 0; s < ns; ++s) {
      const Float mm = (tstep - last_mig_tstep) * migration[s];
      cm[s] = mm > cap ? cap : mm;
    }
    std::vector<std::vector<int32_t>> delta(cells.size(), std::vector<int32_t>(ns, 0));
#pragma omp parallel for schedule(static)
    for (long ci = 0; ci < ncell; ++ci) {
      const Cell& c = cells[ci];
      for (uint32_t b : c.nbr) {
        const Cell& nb = cells[b];
        for (size_t s = 0; s < ns; ++s)
          delta[ci][s] += (int32_t)(cm[s] * (nb.n[s] - c.n[s]));   // Float->int truncation, arrays.h:1408-1411
      }
    }
    uint64_t total = 0;
#pragma omp parallel for reduction(+ : total) schedule(static)
    for (long ci = 0; ci < ncell; ++ci) {
      Cell& c = cells[ci];
      for (size_t s = 0; s < ns; ++s) {
        c.n[s] += delta[ci][s];
        total += (uint64_t)std::abs(delta[ci][s]);
      }
    }
    last_mig_tstep = tstep;
    return total / 2;
  }

  // models/ecolab_model.cc:139-159 ModelData::condense
  void condenseModel(const std::vector<char>& mask) {
    const size_t ns = nsp();
    std::vector<uint32_t> map(ns);   // enumerate(mask), arrays.h:2403-2414
    uint32_t k = 0;
    for (size_t i = 0; i < ns; ++i) { map[i] = k; if (mask[i]) ++k; }
    auto packF = [&](std::vector<Float>& a) {
      std::vector<Float> r; for (size_t i = 0; i < a.size(); ++i) if (mask[i]) r.push_back(a[i]); a.swap(r);
    };
    packF(create);
    { std::vector<int32_t> r; for (size_t i = 0; i < ns; ++i) if (mask[i]) r.push_back(species[i]); species.swap(r); }
    packF(repro_rate); packF(mutation); packF(migration); packF(diag);
    std::vector<Float> v2; std::vector<uint32_t> r2, c2;
    for (size_t j = 0; j < row.size(); ++j)
      if (mask[row[j]] && mask[col[j]]) { v2.push_back(val[j]); r2.push_back(map[row[j]]); c2.push_back(map[col[j]]); }
    val.swap(v2); row.swap(r2); col.swap(c2);
    computeODiagIdx();
  }

  // models/ecolab_model.cc:171-196 SpatialModel::condense, :161-169 PanmicticModel::condense,
  // :124-137 EcolabPoint::condense
  unsigned condense() {
    const size_t ns = panmictic ? cells[0].n.size() : nsp();
    std::vector<char> mask(ns, 0);
    for (auto& c : cells)
      for (size_t s = 0; s < c.n.size() && s < ns; ++s) if (c.n[s]) mask[s] = 1;
    size_t ntrue = 0; for (char m : mask) ntrue += m;
    if (ntrue == ns) return 0;
    condenseModel(mask);
    for (auto& c : cells) {
      std::vector<int32_t> r;
      for (size_t s = 0; s < c.n.size(); ++s) if (mask[s] && r.size() < ntrue) r.push_back(c.n[s]);
      c.n.swap(r);
    }
    return (unsigned)(ns - ntrue);
  }

  // src/arrays.cc:85-102 lgspread / gspread
  static void gspread(std::vector<Float>& a, const std::vector<Float>& s) {
    std::vector<Float> g(a.size());
    for (auto& x : g) x = g_array_grand.rand();
    for (size_t i = 0; i < a.size(); ++i) a[i] += s[i] * g[i];
  }
  static void lgspread(std::vector<Float>& a, const std::vector<Float>& s) {
    std::vector<Float> g(a.size());
    for (auto& x : g) x = g_array_grand.rand();
    for (size_t i = 0; i < a.size(); ++i)
      a[i] = (a[i] >= 0 ? 1 : -1) * std::exp(std::log(std::fabs(a[i])) + s[i] * g[i]);
  }

  // models/ecolab_model.cc:327-359 do_row_or_col
  static void do_row_or_col(std::vector<Float>& tmp, Float range, Float minval, Float gdist, Float gen_bias) {
    double r = (2.0 * ::rand()) / RAND_MAX - 1 + gen_bias;
    r = (r > 0) ? r / (1 + gen_bias) : r / (1 - gen_bias);
    size_t ntrue;
    if (r != 0) {
      const double inv = 1 / std::fabs(r);
      ntrue = inv >= 2147483648.0 ? tmp.size() : (size_t)(unsigned)((int)inv - 1);
    } else
      ntrue = tmp.size();
    if (ntrue > tmp.size()) ntrue = tmp.size();
    if (r > 0)
      for (size_t j = 0; j < ntrue; ++j) {
        const int pos = (int)((tmp.size() - 1) * ((double)::rand() / RAND_MAX) + .5);
        if (tmp[pos] == 0.0) tmp[pos] = range * ((double)::rand() / RAND_MAX) + minval;
      }
    else
      for (size_t j = 0; j < ntrue; ++j) {
        const int pos = (int)((tmp.size() - 1) * ((double)::rand() / RAND_MAX) + .5);
        tmp[pos] = 0;
      }
    std::vector<Float> diff(tmp.size());
    for (size_t i = 0; i < tmp.size(); ++i) diff[i] = tmp[i] != 0.0 ? range * gdist : 0.0;
    gspread(tmp, diff);
  }

  // models/ecolab_model.cc:362-470 ModelData::mutate
  void mutateModel(const std::vector<uint32_t>& new_sp) {
    const size_t m = new_sp.size();
    if (!m) return;
    if (sp_cntr == -1) {
      int mx = INT_MIN; for (int s : species) mx = std::max(mx, s);
      sp_cntr = mx + 1;
    }
    std::vector<Float> gdist(m);
    for (auto& g : gdist) g = -std::log(urand());                 // fillprand, src/arrays.cc:48-52
    for (size_t k = 0; k < m; ++k) gdist[k] *= mutation[new_sp[k]];

    std::vector<Float> nr(m), s(m);
    for (size_t k = 0; k < m; ++k) nr[k] = repro_rate[new_sp[k]];
    double range = repro_max - repro_min;
    for (size_t k = 0; k < m; ++k) s[k] = range * gdist[k];
    gspread(nr, s);
    repro_rate.insert(repro_rate.end(), nr.begin(), nr.end());

    std::vector<Float> nm(m);
    for (size_t k = 0; k < m; ++k) nm[k] = mutation[new_sp[k]];
    if (!fixMutation) lgspread(nm, gdist);
    for (size_t k = 0; k < m; ++k) mutation.push_back(nm[k] < mut_max ? nm[k] : mut_max);

    std::vector<Float> ng(m);
    for (size_t k = 0; k < m; ++k) ng[k] = migration[new_sp[k]];
    if (!fixMigration) lgspread(ng, gdist);
    migration.insert(migration.end(), ng.begin(), ng.end());

    std::vector<Float> nd(m);
    for (size_t k = 0; k < m; ++k) nd[k] = diag[new_sp[k]];
    lgspread(nd, gdist);
    for (size_t k = 0; k < m; ++k) {
      const Float mx = -std::fabs(nr[k]) / (0.1 * INT_MAX);
      diag.push_back(nd[k] < mx ? nd[k] : mx);
    }

    range = odiag_max - odiag_min;
    const size_t ns0 = nsp();   // species not yet appended (:411)
    for (size_t i = 0; i < m; ++i) {
      const size_t L = ns0 + i;
      std::vector<Float> t1(L, 0.0), t2(L, 0.0);
      const uint32_t p = new_sp[i];
      for (size_t j = 0; j < row.size(); ++j) if (row[j] == p) t1[col[j]] = val[j];
      t1[p] = diag[p];
      do_row_or_col(t1, range, odiag_min, gdist[i], gen_bias);
      for (size_t j = 0; j < col.size(); ++j) if (col[j] == p) t2[row[j]] = val[j];
      t2[p] = diag[p];
      do_row_or_col(t2, range, odiag_min, gdist[i], gen_bias);

      Float sum = 0;
      std::vector<char> mask(L);
      for (size_t q = 0; q < L; ++q) { mask[q] = (t1[q] + t2[q] > 0); if (mask[q]) sum += t1[q] + t2[q]; }
      sum += diag[i];   // sic: indexed by the mutant loop counter (:437)
      if (sum > 0) {
        int nadj = 0;
        for (size_t q = 0; q < L; ++q) nadj += (mask[q] && t1[q] != 0.0) || (mask[q] && t2[q] != 0.0);
        for (size_t q = 0; q < L; ++q) {
          const bool m1 = mask[q] && t1[q] != 0.0, m2 = mask[q] && t2[q] != 0.0;
          if (m1) t1[q] = t1[q] - sum / nadj;
          if (m2) t2[q] = t2[q] - sum / nadj;
        }
      }
      for (size_t q = 0; q < L; ++q) if (t1[q] != 0.0) { val.push_back(t1[q]); col.push_back((uint32_t)q); row.push_back((uint32_t)L); }
      for (size_t q = 0; q < L; ++q) if (t2[q] != 0.0) { val.push_back(t2[q]); row.push_back((uint32_t)q); col.push_back((uint32_t)L); }
    }
    for (size_t k = 0; k < m; ++k) { create.push_back(0); species.push_back((int)(k * nprocs) + sp_cntr); }
    sp_cntr += (int)(m * nprocs);
  }

  // models/ecolab_model.cc:215-276 SpatialModel::mutate, :205-213 PanmicticModel::mutate,
  // :278-314 EcolabPoint::mutate (host branch)
  void mutate() {
    const size_t ns = nsp();
    std::vector<Float> ms(ns);
    for (size_t s = 0; s < ns; ++s)
      ms[s] = sp_sep * repro_rate[s] * mutation[s] * Float(tstep - last_mut_tstep);
    last_mut_tstep = tstep;
    std::vector<uint32_t> new_sp, cell_ids;
    for (auto& c : cells) {   // new build: ascending cell id (the reference: unordered_set order)
      if (c.n.empty()) continue;
      std::vector<uint32_t> spec(c.n.size());
      uint64_t tot = 0;
      for (size_t s = 0; s < c.n.size(); ++s) {
        spec[s] = (uint32_t)ROUND(ms[s] * c.n[s], c, ST_MUTATE, (uint32_t)s, tstep);
        tot += spec[s];
      }
      if (!tot) continue;
      for (size_t s = 0; s < c.n.size(); ++s) {
        c.n[s] -= (int32_t)spec[s];
        for (uint32_t k = 0; k < spec[s]; ++k) { new_sp.push_back((uint32_t)s); cell_ids.push_back(c.id); }   // gen_index, arrays.h:2477-2486
      }
    }
    mutateModel(new_sp);
    if (new_sp.empty()) return;
    computeODiagIdx();
    for (auto& c : cells)
      for (uint32_t id : cell_ids) c.n.push_back(id == c.id ? 1 : 0);
  }

  // include/sparse_mat.h:88-111 sparse_mat::init_rand + src/arrays.cc:56-83 fill_unique_rand
  void random_interaction(unsigned conn, double sigma) {
    const size_t n = diag.size();
    std::vector<unsigned> rsize(n);
    for (size_t i = 0; i < n; ++i) {
      const Float t = g_array_grand.rand() * sigma + (double)conn;
      rsize[i] = (unsigned)(t > 0 ? t : 0.0);
    }
    row.clear(); col.clear();
    for (size_t i = 0; i < n; ++i) {
      std::vector<int> c(rsize[i]);
      if (c.size() > n / 2) {
        std::vector<double> v(n);
        for (auto& x : v) x = urand();
        std::vector<int> deck(n);
        for (size_t q = 0; q < n; ++q) deck[q] = (int)q;
        std::sort(deck.begin(), deck.end(), [&](int a, int b) { return v[a] < v[b]; });
        for (size_t q = 0; q < c.size(); ++q) c[q] = deck[q];
      } else {
        std::vector<bool> drawn(n);
        for (size_t q = 0; q < c.size();) {
          const size_t j = (size_t)((n * (double)::rand()) / RAND_MAX);
          if (!drawn[j]) { drawn[j] = true; c[q++] = (int)j; }
        }
      }
      for (int cc : c) if (cc != (int)i) { row.push_back((uint32_t)i); col.push_back((uint32_t)cc); }
    }
    val.assign(row.size(), 0);
  }

  // models/ecolab_model.cc:317-323 ModelData::connectivity
  Float connectivity() const {
    if (species.empty()) return 0;
    Float sum = 0;
    for (size_t j = 0; j < val.size(); ++j) sum += (val[j] * val[j]) / (diag[row[j]] * diag[col[j]]);
    return sum / (species.size() * (species.size() - 1));
  }

  // models/ecolab_model.cc:493-503 PanmicticModel::mutantConnectivity: the masked sum runs in entry order (`sum(x, mask)`,
  // include/arrays.h:2303-2330); note the strict `>` on row/col, as in the reference
  Float mutantConnectivity(size_t beforeNsp) const {
    const size_t n = species.size() - beforeNsp;
    if (species.size() <= beforeNsp) return 0;
    Float sum = 0;
    for (size_t j = 0; j < val.size(); ++j)
      if (row[j] > beforeNsp || col[j] > beforeNsp) sum += (val[j] * val[j]) / (diag[row[j]] * diag[col[j]]);
    return sum / (n * (species.size() - 1));
  }

  // models/ecolab_model.cc:506-516 PanmicticModel::extinctionConnectivity
  Float extinctionConnectivity() const {
    const auto& d = cells[0].n;
    size_t n = 0;
    for (int v : d) n += v == 0;
    if (n == 0) return 0;
    Float sum = 0;
    for (size_t j = 0; j < val.size(); ++j)
      if (d[row[j]] == 0 || d[col[j]] == 0) sum += (val[j] * val[j]) / (diag[row[j]] * diag[col[j]]);
    return sum / (n * (species.size() - 1));
  }

  // models/ecolab_model.cc:474-490 PanmicticModel::lifetimes
  std::vector<Float> lifetimes() {
    std::vector<Float> out;
    const auto& d = cells[0].n;
    for (size_t i = 0; i < species.size(); ++i) {
      if (create[i] == 0 && d[i] > 10) create[i] = (Float)tstep;
      else if (create[i] > 0 && d[i] == 0) { out.push_back(tstep - create[i]); create[i] = 0; }
    }
    return out;
  }
};

std::vector<Float>* f64_member(Oracle* o, const std::string& n) {
  if (n == "create") return &o->create;
  if (n == "repro_rate") return &o->repro_rate;
  if (n == "mutation") return &o->mutation;
  if (n == "migration") return &o->migration;
  if (n == "interaction.diag") return &o->diag;
  if (n == "interaction.val") return &o->val;
  return nullptr;
}
std::vector<uint32_t>* u32_member(Oracle* o, const std::string& n) {
  if (n == "interaction.row") return &o->row;
  if (n == "interaction.col") return &o->col;
  return nullptr;
}
Float* scalar_member(Oracle* o, const std::string& n) {
  if (n == "sp_sep") return &o->sp_sep;
  if (n == "mut_max") return &o->mut_max;
  if (n == "repro_min") return &o->repro_min;
  if (n == "repro_max") return &o->repro_max;
  if (n == "odiag_min") return &o->odiag_min;
  if (n == "odiag_max") return &o->odiag_max;
  if (n == "gen_bias") return &o->gen_bias;
  return nullptr;
}

}  // namespace

// ---------------------------------------------------------------------------------
// Flat C API (ctypes) -- used by tests/, smoke() and bench.py's CPU legs only.
// ---------------------------------------------------------------------------------
extern "C" {

void* orc_create() { return new Oracle(); }
void orc_destroy(void* h) { delete (Oracle*)h; }

void orc_set_rng(void* h, int mode, uint32_t seed) { auto o = (Oracle*)h; o->rng_mode = mode; o->seed = seed; }
// PanmicticModel::seed / SpatialModel::seed, models/ecolab_model.h:111,144
void orc_seed(void* h, uint32_t x) { auto o = (Oracle*)h; o->seed = x; for (auto& c : o->cells) c.mt.seed(x); }
void orc_reset_grand() { g_array_grand = GaussRand(); }
void orc_set_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}
int orc_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

int orc_set_scalar(void* h, const char* name, double v) {
  auto o = (Oracle*)h; std::string n(name);
  if (Float* p = scalar_member(o, n)) { *p = v; return 0; }
  if (n == "tstep") { o->tstep = (uint64_t)v; return 0; }
  if (n == "last_mut_tstep") { o->last_mut_tstep = (uint64_t)v; return 0; }
  if (n == "last_mig_tstep") { o->last_mig_tstep = (uint64_t)v; return 0; }
  if (n == "fixMutation") { o->fixMutation = v != 0; return 0; }
  if (n == "fixMigration") { o->fixMigration = v != 0; return 0; }
  if (n == "record_pre") { o->record_pre = v != 0; return 0; }
  if (n == "nprocs") { o->nprocs = (unsigned)v; return 0; }
  return -1;
}
double orc_get_scalar(void* h, const char* name) {
  auto o = (Oracle*)h; std::string n(name);
  if (Float* p = scalar_member(o, n)) return *p;
  if (n == "tstep") return (double)o->tstep;
  if (n == "last_mut_tstep") return (double)o->last_mut_tstep;
  if (n == "last_mig_tstep") return (double)o->last_mig_tstep;
  if (n == "fixMutation") return o->fixMutation;
  if (n == "fixMigration") return o->fixMigration;
  if (n == "nsp") return (double)o->nsp();
  if (n == "nnz") return (double)o->val.size();
  if (n == "ncells") return (double)o->cells.size();
  if (n == "maxNbrs") return (double)o->maxNbrs;
  return NAN;
}
int orc_set_f64(void* h, const char* name, const double* p, size_t n) {
  auto v = f64_member((Oracle*)h, name); if (!v) return -1; v->assign(p, p + n); return 0;
}
long orc_get_f64(void* h, const char* name, double* p, size_t cap) {
  auto v = f64_member((Oracle*)h, name); if (!v) return -1;
  if (p) std::copy(v->begin(), v->begin() + std::min(cap, v->size()), p);
  return (long)v->size();
}
int orc_set_u32(void* h, const char* name, const uint32_t* p, size_t n) {
  auto v = u32_member((Oracle*)h, name); if (!v) return -1; v->assign(p, p + n); return 0;
}
long orc_get_u32(void* h, const char* name, uint32_t* p, size_t cap) {
  auto v = u32_member((Oracle*)h, name); if (!v) return -1;
  if (p) std::copy(v->begin(), v->begin() + std::min(cap, v->size()), p);
  return (long)v->size();
}
void orc_set_species(void* h, const int32_t* p, size_t n) { ((Oracle*)h)->species.assign(p, p + n); }
long orc_get_species(void* h, int32_t* p, size_t cap) {
  auto& v = ((Oracle*)h)->species;
  if (p) std::copy(v.begin(), v.begin() + std::min(cap, v.size()), p);
  return (long)v.size();
}

void orc_set_grid(void* h, size_t nx, size_t ny) { ((Oracle*)h)->setGrid(nx, ny); }
void orc_set_grid_slab(void* h, size_t nx, size_t ny, size_t y0, size_t nyl) { ((Oracle*)h)->setGridSlab(nx, ny, y0, nyl); }
uint64_t orc_migrate_local(void* h) { return ((Oracle*)h)->migrateLocal(); }
void orc_set_ghost(void* h, int bottom, const int32_t* p, size_t n) {
  auto o = (Oracle*)h; (bottom ? o->ghost_bottom : o->ghost_top).assign(p, p + n);
}
void orc_alive_mask(void* h, uint8_t* mask) { ((Oracle*)h)->aliveMask(mask); }
unsigned orc_condense_apply(void* h, const uint8_t* mask) { return ((Oracle*)h)->condenseApply(mask); }
size_t orc_mutate_cells(void* h) { auto o = (Oracle*)h; o->mutateCells(); return o->pending.size(); }
void orc_mutate_records(void* h, uint32_t* out) {
  auto o = (Oracle*)h;
  for (size_t i = 0; i < o->pending.size(); ++i) { out[3 * i] = o->pending[i].cell; out[3 * i + 1] = o->pending[i].slot; out[3 * i + 2] = o->pending[i].count; }
}
void orc_append_species(void* h, const uint32_t* ids, size_t m) { auto o = (Oracle*)h; o->computeODiagIdx(); o->appendSpecies(ids, m); }
// synthetic initial state shared with the engine's eco_fill_density_random (Philox stage 3, t = 0)
void orc_fill_density_random(void* h, int32_t lo, int32_t hi) {
  auto o = (Oracle*)h;
  const size_t ns = o->nsp();
  const uint32_t span = (uint32_t)(hi - lo) + 1u;
  for (auto& c : o->cells) {
    c.n.resize(ns);
    for (size_t s = 0; s < ns; ++s) c.n[s] = lo + (int32_t)(philox_draw(o->seed, 3, (uint32_t)(c.id % o->numX), (uint32_t)(c.id / o->numX), (uint32_t)s, 0) % span);
  }
}
void orc_set_density(void* h, size_t cell, const int32_t* p, size_t n) { ((Oracle*)h)->cells[cell].n.assign(p, p + n); }
long orc_get_density(void* h, size_t cell, int32_t* p, size_t cap) {
  auto& v = ((Oracle*)h)->cells[cell].n;
  if (p) std::copy(v.begin(), v.begin() + std::min(cap, v.size()), p);
  return (long)v.size();
}
// all cells, cell-major [cell id][nsp]
void orc_set_density_all(void* h, const int32_t* p, size_t nsp) {
  auto o = (Oracle*)h;
  for (size_t c = 0; c < o->cells.size(); ++c) o->cells[c].n.assign(p + c * nsp, p + (c + 1) * nsp);
}
void orc_get_density_all(void* h, int32_t* p, size_t nsp) {
  auto o = (Oracle*)h;
  for (size_t c = 0; c < o->cells.size(); ++c) std::copy(o->cells[c].n.begin(), o->cells[c].n.begin() + nsp, p + c * nsp);
}
double orc_get_salt(void* h, size_t cell) { return ((Oracle*)h)->cells[cell].salt; }

void orc_make_consistent(void* h) { ((Oracle*)h)->makeConsistent(); }
void orc_generate(void* h, unsigned niter) { ((Oracle*)h)->generate(niter); }
uint64_t orc_migrate(void* h) { return ((Oracle*)h)->migrate(); }
unsigned orc_condense(void* h) { return ((Oracle*)h)->condense(); }
void orc_mutate(void* h) { ((Oracle*)h)->mutate(); }
void orc_mutate_model(void* h, const uint32_t* parents, size_t m) {
  ((Oracle*)h)->mutateModel(std::vector<uint32_t>(parents, parents + m));
}
void orc_random_interaction(void* h, unsigned conn, double sigma) { ((Oracle*)h)->random_interaction(conn, sigma); }
double orc_connectivity(void* h) { return ((Oracle*)h)->connectivity(); }
double orc_mutant_connectivity(void* h, size_t before) { return ((Oracle*)h)->mutantConnectivity(before); }
double orc_extinction_connectivity(void* h) { return ((Oracle*)h)->extinctionConnectivity(); }
long orc_lifetimes(void* h, double* out, size_t cap) {
  auto v = ((Oracle*)h)->lifetimes();
  std::copy(v.begin(), v.begin() + std::min(cap, v.size()), out);
  return (long)v.size();
}
void orc_nsp_per_cell(void* h, uint32_t* out) {   // EcolabPoint::nsp, models/ecolab_model.cc:114-122
  auto o = (Oracle*)h;
  for (size_t c = 0; c < o->cells.size(); ++c) {
    uint32_t k = 0; for (int v : o->cells[c].n) k += v != 0; out[c] = k;
  }
}
long orc_get_pre(void* h, double* p, size_t cap) {
  auto& v = ((Oracle*)h)->pre;
  if (p) std::copy(v.begin(), v.begin() + std::min(cap, v.size()), p);
  return (long)v.size();
}
uint32_t orc_philox(uint32_t seed, uint32_t stage, uint32_t x, uint32_t y, uint32_t slot, uint64_t t) {
  return philox_draw(seed, stage, x, y, slot, t);
}
void orc_philox_raw(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {   // known-answer tests
  uint32_t c[4] = {ctr[0], ctr[1], ctr[2], ctr[3]};
  philox_block(c, key[0], key[1]);
  for (int i = 0; i < 4; ++i) out[i] = c[i];
}
int orc_round(double x, uint32_t u) {   // ROUND with an injected draw (unit tests)
  const Float maxInt = Float(INT_MAX - 1);
  if (x <= 0) return 0;
  if (x >= maxInt) return (int)maxInt;
  Float ip; const Float frac = std::fabs(std::modf(x, &ip));
  return frac * Float(4294967295u) > Float(u) ? (int)x + 1 : (int)x;
}

}  // extern "C"
