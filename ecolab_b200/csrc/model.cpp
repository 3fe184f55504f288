// model.cpp -- host side of the engine: the ModelData / PanmicticModel /
// SpatialModel state machine of the reference (models/ecolab_model.{h,cc}) kept
// in C++ above the CUDA device store, and the C ABI declared in
// include/ecolab_b200.h.  Species-level bookkeeping (condense of the parameter
// arrays, parameter mutation) is O(nsp + nnz) serial host work exactly as in the
// reference; everything O(cells x species) runs in engine.cu on the GPU.
#include <algorithm>
#include <iterator>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/ecolab_b200.h"
#include "engine.h"

namespace {

thread_local std::string g_err;

// ---- the reference's process-wide generators (src/arrays.cc:24-25) -----------------
// array_urand: libc rand()/RAND_MAX (src/random.cc:57); array_grand: running-CLT
// generator (src/random.cc:59-64).  Kept semantically identical so that
// `array_urand.seed(n)` in the driver scripts means the same thing.
//
// rand() here is glibc's: the additive feedback generator r[i] = r[i-3] + r[i-31] (TYPE_3 of random_r, seeded by the
// Lehmer sequence 16807*x mod 2^31-1, first 310 outputs discarded, low bit dropped), restated so that the model OWNS the
// state.  Two reasons: ModelData::mutate draws a gaussian for every element of two dense temporaries per mutant
// (do_row_or_col, :351-358) -- millions of calls per mutate at 1000 species -- and libc's rand() takes a lock per call
// (29 ns vs 2 ns here); and an owned state goes into the checkpoint, so a restarted run draws the same parameters.
// The tests run it side by side with libc rand() (the oracle's) from the same seeds: identical streams.
constexpr int kRandMax = 2147483647;
struct GlibcRand {
  int32_t r[31];
  int32_t f = 3, b = 0;
  GlibcRand() { seed(1); }   // an unseeded libc generator behaves like srand(1)
  void seed(unsigned s) {
    if (s == 0) s = 1;
    r[0] = (int32_t)s;
    int32_t word = (int32_t)s;
    for (int i = 1; i < 31; ++i) {
      const long hi = word / 127773, lo = word % 127773;
      word = (int32_t)(16807 * lo - 2836 * hi);
      if (word < 0) word += 2147483647;
      r[i] = word;
    }
    f = 3; b = 0;
    for (int k = 0; k < 310; ++k) next();
  }
  int next() {
    const uint32_t val = (uint32_t)r[f] + (uint32_t)r[b];
    r[f] = (int32_t)val;
    if (++f >= 31) { f = 0; ++b; } else if (++b >= 31) b = 0;
    return (int)(val >> 1);
  }
} g_urand;
inline double urand01() { return ((double)g_urand.next()) / kRandMax; }
struct RunningClt {
  double sum = 0;
  unsigned n = 0;
  double next() {
    sum += urand01();
    ++n;
    return (sum - .5 * n) * std::sqrt(12. / n);
  }
  void skip() {   // advance the state past a draw nobody looks at (no sqrt, no division by n)
    sum += urand01();
    ++n;
  }
} g_grand;

// host copy of the Philox4x32-10 contract (k_common.cuh), only for the `salt` getter:
// block = philox(ctr = {(x>>2) | stage<<30, y, slot, t_lo}, key = {seed, t_hi}); draw = block[x & 3]
uint32_t philox_host(uint32_t seed, uint32_t stage, uint32_t x, uint32_t y, uint32_t slot, uint64_t t) {
  uint32_t c0 = (x >> 2) | (stage << 30), c1 = y, c2 = slot, c3 = (uint32_t)t, k0 = seed, k1 = (uint32_t)(t >> 32);
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  const uint32_t o[4] = {c0, c1, c2, c3};
  return o[x & 3];
}

void add_noise(std::vector<double>& a, const std::vector<double>& scale) {          // gspread, src/arrays.cc:94-100
  std::vector<double> z(a.size());
  for (double& v : z) v = g_grand.next();
  for (size_t i = 0; i < a.size(); ++i) a[i] += scale[i] * z[i];
}
void mul_lognoise(std::vector<double>& a, const std::vector<double>& scale) {       // lgspread, src/arrays.cc:85-92
  std::vector<double> z(a.size());
  for (double& v : z) v = g_grand.next();
  for (size_t i = 0; i < a.size(); ++i) {
    const double sgn = a[i] >= 0 ? 1.0 : -1.0;
    a[i] = sgn * std::exp(std::log(std::fabs(a[i])) + scale[i] * z[i]);
  }
}

}  // namespace

struct eco_model {
  // ---- ModelData (models/ecolab_model.h:39-68) ----
  std::vector<int32_t> species;
  std::vector<double> create, repro_rate, mutation, migration, diag, val;
  std::vector<uint32_t> row, col;
  uint64_t tstep = 0, last_mut_tstep = 0, last_mig_tstep = 0;
  double sp_sep = 0.1, mut_max = 0, repro_min = 0, repro_max = 1, odiag_min = 0, odiag_max = 1, gen_bias = 0;
  bool fixMutation = false, fixMigration = false;
  // ---- SpatialModel (models/ecolab_model.h:128-151) ----
  bool panmictic = true;
  size_t numX = 1, numY = 1, maxNbrs = 0;
  size_t y0 = 0, nyl = 1;      // owned slab
  unsigned nprocs = 1;
  // ---- engine ----
  eco::Engine eng;
  bool beta_dirty = true;
  uint32_t seed = 0;
  uint64_t salt_t = 0;
  int sp_cntr = -1;   // the reference's function-static species label counter (models/ecolab_model.cc:366)
  std::vector<eco::MutRecord> pending;   // records of the last eco_mutate_cells
  bool host_begun = false;               // eco_step_host_begin ran; eco_step_host finishes the step

  explicit eco_model(int dev) : eng(dev) { eng.set_grid(1, 1, 0, 1, 0); }

  size_t nsp() const { return species.size(); }
  size_t ncells_local() const { return numX * nyl; }

  void grow_engine(size_t n) {
    if (n > eng.nsp()) { eng.resize_species((uint32_t)n); }
  }

  // ModelData::makeConsistent :705-716 + SpatialModel::makeConsistent :725-735
  void makeConsistent() {
    size_t n = std::max(nsp(), (size_t)eng.nsp());
    if (species.empty()) {
      species.resize(n);
      for (size_t i = 0; i < n; ++i) species[i] = (int32_t)i;
    }
    if (create.empty()) create.assign(nsp(), 0);
    if (mutation.empty()) mutation.assign(nsp(), 0);
    if (migration.empty()) migration.assign(nsp(), 0);
    grow_engine(nsp());
    beta_dirty = true;   // computeODiagIdx :718-723 happens in upload()
  }

  void check_sizes(bool need_migration) {
    const size_t n = nsp();
    auto bad = [&](const char* what, size_t have) {
      char b[256];
      snprintf(b, sizeof b, "inconsistent model: %s has %zu entries but there are %zu species", what, have, n);
      throw std::runtime_error(b);
    };
    if (eng.nsp() > n) bad("cell density", eng.nsp());
    grow_engine(n);
    // migrate reads only `migration` (:626-629): models/migration_test.py sets nothing else before calling it
    if (need_migration) {
      if (migration.size() != n) bad("migration", migration.size());
      return;
    }
    if (repro_rate.size() != n) bad("repro_rate", repro_rate.size());
    if (diag.size() != n) bad("interaction.diag", diag.size());
    if (row.size() != val.size() || col.size() != val.size())
      throw std::runtime_error("inconsistent model: interaction.row/col/val differ in length");
  }

  // hand beta to the engine as CSR; term order inside a row = COO order (computeODiagIdx, :718-723)
  void upload() {
    if (!beta_dirty) return;
    const uint32_t n = (uint32_t)nsp();
    eco::BetaCsr b;
    b.nsp = n;
    b.rowptr.assign(n + 1, 0);
    for (size_t j = 0; j < row.size(); ++j) {
      if (row[j] >= n || col[j] >= n) throw std::runtime_error("interaction.row/col index out of range");
      ++b.rowptr[row[j] + 1];
    }
    for (uint32_t i = 0; i < n; ++i) b.rowptr[i + 1] += b.rowptr[i];
    std::vector<uint32_t> pos(b.rowptr.begin(), b.rowptr.end() - 1);
    b.val.resize(row.size());
    b.col.resize(row.size());
    for (size_t j = 0; j < row.size(); ++j) {   // stable: ascending j inside a row
      const uint32_t k = pos[row[j]]++;
      b.val[k] = val[j];
      b.col[k] = col[j];
    }
    b.diag = diag;
    b.rr = repro_rate;
    eng.upload_beta(std::move(b));
    beta_dirty = false;
  }

  // SpatialModel::setGrid :580-605
  void setGrid(size_t nx, size_t ny, size_t y0_, size_t nyl_) {
    if (!nx || !ny) throw std::runtime_error("setGrid: empty grid");
    if (nx * ny > 0xffffffffull) throw std::runtime_error("setGrid: more than 2^32 cells");
    if (y0_ + nyl_ > ny) throw std::runtime_error("setGrid: slab outside the grid");
    panmictic = false;
    numX = nx; numY = ny; y0 = y0_; nyl = nyl_;
    eng.set_grid((uint32_t)nx, (uint32_t)ny, (uint32_t)y0, (uint32_t)nyl, (uint32_t)nsp());
    const size_t nb = std::min<size_t>(nx - 1, 2) + std::min<size_t>(ny - 1, 2);
    maxNbrs = std::max(maxNbrs, nb);   // never shrinks (:603-604)
  }

  size_t local_cell(size_t cell) const {
    if (cell >= numX * numY) throw std::runtime_error("cell id outside the grid");
    const size_t y = cell / numX;
    if (y < y0 || y >= y0 + nyl) throw std::runtime_error("cell is not owned by this process");
    return cell - y0 * numX;
  }

  void generate(unsigned niter) {
    if (tstep == 0) makeConsistent();
    check_sizes(false);
    upload();
    eng.generate(niter, tstep, seed);
    tstep += niter;
  }

  uint64_t migrate_local(bool wait = true) {
    check_sizes(true);
    const size_t n = nsp();
    std::vector<double> cm(n);
    const double cap = 1.0 / maxNbrs;   // :628
    for (size_t s = 0; s < n; ++s) {
      const double mm = (tstep - last_mig_tstep) * migration[s];   // :626
      cm[s] = mm > cap ? cap : mm;                                  // :629
    }
    salt_t = tstep;
    const uint64_t total = n ? eng.migrate(cm.data(), wait) : 0;
    last_mig_tstep = tstep;
    return total;
  }

  // generate(niter) then migrate() fed from / drained to host buffers (cell-major [cell][nsp], ascending cell id)
  bool partitioned() const { return y0 != 0 || nyl != numY; }
  void step_host_check(size_t ncells, size_t ns) {
    if (panmictic) throw std::runtime_error("step_host: needs a spatial grid");
    if (ncells != ncells_local()) throw std::runtime_error("step_host: cell count differs from the owned grid");
    if (tstep == 0) makeConsistent();
    grow_engine(std::max(ns, nsp()));
    check_sizes(false);
    check_sizes(true);
    if (ns != nsp()) throw std::runtime_error("step_host: species count differs from the model");
    upload();
  }
  void step_host_begin(const int32_t* in, size_t ncells, size_t ns, unsigned niter) {
    step_host_check(ncells, ns);
    if (nsp()) eng.step_host_begin(in, niter, tstep, seed);
    host_begun = true;
  }
  uint64_t step_host(const int32_t* in, int32_t* out, size_t ncells, size_t ns, unsigned niter) {
    const bool part = partitioned();
    if (part && !host_begun) throw std::runtime_error("step_host: on a partitioned grid call eco_step_host_begin and exchange the halo first");
    if (!part) step_host_check(ncells, ns);
    host_begun = false;
    const size_t n = nsp();
    std::vector<double> cm(n);
    const double cap = 1.0 / maxNbrs;                                   // :628
    const uint64_t t_after = tstep + niter;
    for (size_t s = 0; s < n; ++s) {
      const double mm = (t_after - last_mig_tstep) * migration[s];      // :626 (migrate runs after tstep += niter)
      cm[s] = mm > cap ? cap : mm;                                      // :629
    }
    const uint64_t total = n ? eng.step_host(in, out, niter, tstep, seed, cm.data(), part) : 0;
    tstep = t_after;
    salt_t = tstep;
    last_mig_tstep = tstep;
    return total;
  }

  // ModelData::condense :139-159
  void condenseModel(const std::vector<uint8_t>& mask) {
    const size_t n = nsp();
    std::vector<uint32_t> map(n);
    uint32_t k = 0;
    for (size_t i = 0; i < n; ++i) { map[i] = k; k += mask[i] ? 1 : 0; }
    auto squeeze = [&](auto& a) {
      size_t w = 0;
      for (size_t i = 0; i < a.size() && i < n; ++i) if (mask[i]) a[w++] = a[i];
      a.resize(w);
    };
    squeeze(create); squeeze(species); squeeze(repro_rate); squeeze(mutation); squeeze(migration); squeeze(diag);
    size_t w = 0;
    for (size_t j = 0; j < val.size(); ++j)
      if (mask[row[j]] && mask[col[j]]) { val[w] = val[j]; row[w] = map[row[j]]; col[w] = map[col[j]]; ++w; }
    val.resize(w); row.resize(w); col.resize(w);
    beta_dirty = true;
  }

  unsigned condense_apply(const std::vector<uint8_t>& mask) {
    const size_t n = nsp();
    size_t ntrue = 0;
    for (uint8_t b : mask) ntrue += b ? 1 : 0;
    if (ntrue == n) return 0;
    condenseModel(mask);
    eng.compact_species(mask.data(), (uint32_t)ntrue);
    return (unsigned)(n - ntrue);
  }

  unsigned condense() {
    if (panmictic && species.empty()) makeConsistent();
    const size_t n = nsp();
    if (eng.nsp() != n) check_sizes(false);
    std::vector<uint8_t> mask(n, 0);
    if (n) eng.alive_mask(mask.data());
    return condense_apply(mask);
  }

  // do_row_or_col, models/ecolab_model.cc:327-359, on the first L elements of the dense temporary `t`; appends to `nz` the
  // (ascending) indices that held a link when the gaussians were applied -- the only elements later steps can find non-zero
  void perturb_links(std::vector<double>& t, size_t L, double range, double minval, double gd, std::vector<uint32_t>& nz) {
    double r = (2.0 * g_urand.next()) / kRandMax - 1 + gen_bias;
    r = r > 0 ? r / (1 + gen_bias) : r / (1 - gen_bias);
    size_t cnt = L;
    if (r != 0) {
      const double inv = 1 / std::fabs(r);
      if (inv < 2147483648.0) cnt = (size_t)(unsigned)((int)inv - 1);
    }
    cnt = std::min(cnt, L);
    const double span = (double)(L - 1);
    for (size_t j = 0; j < cnt; ++j) {
      const int pos = (int)(span * ((double)g_urand.next() / kRandMax) + .5);
      if (r > 0) {
        if (t[pos] == 0.0) t[pos] = range * ((double)g_urand.next() / kRandMax) + minval;
      } else
        t[pos] = 0;
    }
    // gspread with scale = (t != 0 ? range*gd : 0) (:351-358): the reference draws a gaussian for EVERY element of the dense
    // temporary; the draws of the zero elements are multiplied by 0, so only the generator's state has to move past them
    // (t + 0*z == t exactly: z is finite).  This loop is where ModelData::mutate spends its time -- two temporaries of
    // nsp elements per mutant, a handful of them non-zero.  The uniforms are drawn a block at a time (the divisions by
    // RAND_MAX pipeline instead of sitting between two steps of the running sum); the sum itself stays one addition after
    // the other, in draw order: it is not associative.
    const double s = range * gd;
    double u[64];
    for (size_t i0 = 0; i0 < L; i0 += 64) {
      const size_t nb = std::min<size_t>(64, L - i0);
      for (size_t k = 0; k < nb; ++k) u[k] = urand01();
      double sum = g_grand.sum;
      unsigned n = g_grand.n;
      for (size_t k = 0; k < nb; ++k) {
        sum += u[k];
        ++n;
        double& v = t[i0 + k];
        if (v != 0.0) {
          v += s * ((sum - .5 * n) * std::sqrt(12. / n));
          nz.push_back((uint32_t)(i0 + k));
        }
      }
      g_grand.sum = sum;
      g_grand.n = n;
    }
  }

  // ModelData::mutate :362-470 -- one new species per entry of `parents`
  void mutateModel(const std::vector<uint32_t>& parents) {
    const size_t m = parents.size();
    if (!m) return;
    const size_t n0 = nsp();
    for (uint32_t p : parents) if (p >= n0) throw std::runtime_error("mutate: parent slot out of range");
    for (size_t j = 0; j < row.size(); ++j)   // before anything is modified
      if (row[j] >= n0 || col[j] >= n0) throw std::runtime_error("mutate: interaction.row/col index out of range");
    if (sp_cntr == -1) sp_cntr = *std::max_element(species.begin(), species.end()) + 1;

    std::vector<double> gd(m);                                   // genetic distances (:371-373)
    for (double& g : gd) g = -std::log(urand01());
    for (size_t k = 0; k < m; ++k) gd[k] *= mutation[parents[k]];

    std::vector<double> r_new(m), sc(m);
    for (size_t k = 0; k < m; ++k) { r_new[k] = repro_rate[parents[k]]; sc[k] = (repro_max - repro_min) * gd[k]; }
    add_noise(r_new, sc);                                        // :380-383
    repro_rate.insert(repro_rate.end(), r_new.begin(), r_new.end());

    std::vector<double> mu(m);
    for (size_t k = 0; k < m; ++k) mu[k] = mutation[parents[k]];
    if (!fixMutation) mul_lognoise(mu, gd);
    for (size_t k = 0; k < m; ++k) mutation.push_back(mu[k] < mut_max ? mu[k] : mut_max);   // :389

    std::vector<double> mg(m);
    for (size_t k = 0; k < m; ++k) mg[k] = migration[parents[k]];
    if (!fixMigration) mul_lognoise(mg, gd);
    migration.insert(migration.end(), mg.begin(), mg.end());

    std::vector<double> dg(m);
    for (size_t k = 0; k < m; ++k) dg[k] = diag[parents[k]];
    mul_lognoise(dg, gd);
    for (size_t k = 0; k < m; ++k) {
      const double lim = -std::fabs(r_new[k]) / (0.1 * INT_MAX);   // :402
      diag.push_back(dg[k] < lim ? dg[k] : lim);
    }

    const double range = odiag_max - odiag_min;
    // entries of every row / column in COO order (the reference scans the whole COO per mutant, :414-423; the lists give
    // the same scatter order -- a later duplicate overwrites an earlier one -- in O(row length)); entries appended by an
    // earlier mutant of this batch are seen by a later one, as in the reference.
    // The reference's dense temporaries of nsp+i elements per mutant are kept (the gaussians are drawn per ELEMENT), but
    // everything after the draws walks only the elements that can be non-zero: the ascending union of the two index lists
    // perturb_links returns -- the same elements, in the same order, as the reference's dense loops meet them.
    // The lists of the existing entries are two counting sorts into flat arrays (stable: ascending j inside a row / column);
    // entries appended during this batch go to small per-species lists behind them -- only parents (< n0) are ever looked up.
    const size_t nnz0 = row.size();
    std::vector<uint32_t> rstart(n0 + 1, 0), cstart(n0 + 1, 0), ridx(nnz0), cidx(nnz0);
    for (size_t j = 0; j < nnz0; ++j) { ++rstart[row[j] + 1]; ++cstart[col[j] + 1]; }
    for (size_t i = 0; i < n0; ++i) { rstart[i + 1] += rstart[i]; cstart[i + 1] += cstart[i]; }
    {
      std::vector<uint32_t> rp(rstart.begin(), rstart.end() - 1), cp(cstart.begin(), cstart.end() - 1);
      for (size_t j = 0; j < nnz0; ++j) { ridx[rp[row[j]]++] = (uint32_t)j; cidx[cp[col[j]]++] = (uint32_t)j; }
    }
    std::vector<std::vector<uint32_t>> xrow(n0), xcol(n0);
    // f(j) for every entry j of row / column p, in COO order
    auto each_in_row = [&](uint32_t p, auto&& f) {
      for (uint32_t k = rstart[p]; k < rstart[p + 1]; ++k) f(ridx[k]);
      for (uint32_t j : xrow[p]) f(j);
    };
    auto each_in_col = [&](uint32_t p, auto&& f) {
      for (uint32_t k = cstart[p]; k < cstart[p + 1]; ++k) f(cidx[k]);
      for (uint32_t j : xcol[p]) f(j);
    };
    std::vector<double> rw(n0 + m, 0.0), cl(n0 + m, 0.0);
    std::vector<uint32_t> nzr, nzc, both;
    for (size_t i = 0; i < m; ++i) {
      const size_t L = n0 + i;   // index of the new species; length of the dense temporaries (:411)
      const uint32_t p = parents[i];
      nzr.clear(); nzc.clear(); both.clear();
      each_in_row(p, [&](uint32_t j) { if (col[j] < L) rw[col[j]] = val[j]; });
      rw[p] = diag[p];
      perturb_links(rw, L, range, odiag_min, gd[i], nzr);
      each_in_col(p, [&](uint32_t j) { if (row[j] < L) cl[row[j]] = val[j]; });
      cl[p] = diag[p];
      perturb_links(cl, L, range, odiag_min, gd[i], nzc);
      std::set_union(nzr.begin(), nzr.end(), nzc.begin(), nzc.end(), std::back_inserter(both));

      // keep n.beta.n < 0: shave the positive part (:434-446); note diag is indexed by i, as in the reference (:437)
      double s = 0;
      int nadj = 0;
      for (uint32_t q : both)
        if (rw[q] + cl[q] > 0) {
          s += rw[q] + cl[q];
          nadj += (rw[q] != 0.0 || cl[q] != 0.0);
        }
      s += diag[i];
      if (s > 0) {
        for (uint32_t q : both) {
          const bool pos = rw[q] + cl[q] > 0;
          const bool a = pos && rw[q] != 0.0, b = pos && cl[q] != 0.0;
          if (a) rw[q] = rw[q] - s / nadj;
          if (b) cl[q] = cl[q] - s / nadj;
        }
      }
      for (uint32_t q : nzr)
        if (rw[q] != 0.0) {
          if (q < n0) xcol[q].push_back((uint32_t)val.size());     // entry (L, q): a new element of column q
          val.push_back(rw[q]); row.push_back((uint32_t)L); col.push_back((uint32_t)q);
        }
      for (uint32_t q : nzc)
        if (cl[q] != 0.0) {
          if (q < n0) xrow[q].push_back((uint32_t)val.size());     // entry (q, L): a new element of row q
          val.push_back(cl[q]); row.push_back((uint32_t)q); col.push_back((uint32_t)L);
        }
      // back to all-zero for the next mutant: everything that was written is in the scatter lists or in nzr / nzc
      each_in_row(p, [&](uint32_t j) { if (col[j] < rw.size()) rw[col[j]] = 0.0; });
      each_in_col(p, [&](uint32_t j) { if (row[j] < cl.size()) cl[row[j]] = 0.0; });
      rw[p] = 0.0; cl[p] = 0.0;
      for (uint32_t q : nzr) rw[q] = 0.0;
      for (uint32_t q : nzc) cl[q] = 0.0;
    }
    for (size_t k = 0; k < m; ++k) { create.push_back(0); species.push_back((int32_t)(k * nprocs) + sp_cntr); }
    sp_cntr += (int)(m * nprocs);
    beta_dirty = true;
  }

  void mutate_cells(std::vector<eco::MutRecord>& rec) {
    check_sizes(false);
    const size_t n = nsp();
    if (mutation.size() != n || migration.size() != n || create.size() != n)
      throw std::runtime_error("inconsistent model: mutation/migration/create size differs from species (call makeConsistent)");
    std::vector<double> ms(n);
    const double dt = (double)(tstep - last_mut_tstep);
    for (size_t s = 0; s < n; ++s) ms[s] = sp_sep * repro_rate[s] * mutation[s] * dt;   // :207 / :218
    last_mut_tstep = tstep;
    eng.mutate_cells(ms.data(), tstep, seed, rec);
  }

  void mutate() {
    std::vector<eco::MutRecord> rec;
    mutate_cells(rec);
    std::vector<uint32_t> parents, cells;
    for (const auto& r : rec)
      for (uint32_t k = 0; k < r.count; ++k) { parents.push_back(r.slot); cells.push_back(r.cell); }   // gen_index
    mutateModel(parents);
    if (parents.empty()) return;
    eng.append_species(cells.data(), cells.size());
  }

  // sparse_mat::init_rand, include/sparse_mat.h:88-111; fill_unique_rand, src/arrays.cc:56-83
  void random_interaction(unsigned conn, double sigma) {
    const size_t n = diag.size();
    if (conn >= n && n) throw std::runtime_error("random_interaction: conn must be smaller than the number of species");
    std::vector<unsigned> rsize(n);
    for (size_t i = 0; i < n; ++i) {
      const double t = g_grand.next() * sigma + (double)conn;
      rsize[i] = (unsigned)(t > 0 ? t : 0.0);
    }
    row.clear(); col.clear();
    for (size_t i = 0; i < n; ++i) {
      std::vector<int> pick(std::min<size_t>(rsize[i], n));
      if (pick.size() > n / 2) {
        std::vector<double> key(n);
        for (double& v : key) v = urand01();
        std::vector<int> deck(n);
        for (size_t q = 0; q < n; ++q) deck[q] = (int)q;
        std::sort(deck.begin(), deck.end(), [&](int a, int b) { return key[a] < key[b]; });
        std::copy(deck.begin(), deck.begin() + pick.size(), pick.begin());
      } else {
        std::vector<bool> seen(n);
        for (size_t q = 0; q < pick.size();) {
          const size_t j = (size_t)((n * (double)g_urand.next()) / kRandMax);
          if (j < n && !seen[j]) { seen[j] = true; pick[q++] = (int)j; }
        }
      }
      for (int c : pick) if (c != (int)i) { row.push_back((uint32_t)i); col.push_back((uint32_t)c); }
    }
    val.assign(row.size(), 0.0);
    beta_dirty = true;
  }

  double connectivity() const {   // :317-323
    if (species.empty()) return 0;
    double s = 0;
    for (size_t j = 0; j < val.size(); ++j) s += (val[j] * val[j]) / (diag[row[j]] * diag[col[j]]);
    return s / (species.size() * (species.size() - 1));
  }
};

// ---------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------
#define ECO_TRY try {
#define ECO_CATCH                                   \
  }                                                 \
  catch (const std::exception& e) {                 \
    g_err = e.what();                               \
    return 1;                                       \
  }                                                 \
  catch (...) {                                     \
    g_err = "unknown error";                        \
    return 1;                                       \
  }                                                 \
  return 0;

namespace {
std::vector<double>* f64_member(eco_model* m, const std::string& n) {
  if (n == "create") return &m->create;
  if (n == "repro_rate") return &m->repro_rate;
  if (n == "mutation") return &m->mutation;
  if (n == "migration") return &m->migration;
  if (n == "interaction.diag") return &m->diag;
  if (n == "interaction.val") return &m->val;
  return nullptr;
}
double* scalar_member(eco_model* m, const std::string& n) {
  if (n == "sp_sep") return &m->sp_sep;
  if (n == "mut_max") return &m->mut_max;
  if (n == "repro_min") return &m->repro_min;
  if (n == "repro_max") return &m->repro_max;
  if (n == "odiag_min") return &m->odiag_min;
  if (n == "odiag_max") return &m->odiag_max;
  if (n == "gen_bias") return &m->gen_bias;
  return nullptr;
}
template <class T>
void copy_out(const std::vector<T>& v, T* out, size_t cap, size_t* n) {
  if (n) *n = v.size();
  if (out) std::copy(v.begin(), v.begin() + std::min(cap, v.size()), out);
}
void need(eco_model* m) {
  if (!m) throw std::runtime_error("null model handle");
}
}  // namespace

extern "C" {

const char* eco_last_error(void) { return g_err.c_str(); }
const char* eco_version(void) { return "ecolab_b200 0.1 (sm_100a)"; }

int eco_create(eco_model** out, int dev) {
  ECO_TRY
  if (!out) throw std::runtime_error("null out pointer");
  *out = new eco_model(dev);
  ECO_CATCH
}
int eco_destroy(eco_model* m) {
  ECO_TRY
  delete m;
  ECO_CATCH
}
int eco_set_stream(eco_model* m, void* s) {
  ECO_TRY
  need(m);
  m->eng.set_stream(s);
  ECO_CATCH
}
int eco_sync(eco_model* m) {
  ECO_TRY
  need(m);
  m->eng.sync();
  ECO_CATCH
}

int eco_set_f64(eco_model* m, const char* name, const double* v, size_t n) {
  ECO_TRY
  need(m);
  auto* a = f64_member(m, name);
  if (!a) throw std::runtime_error(std::string("unknown f64 member: ") + name);
  a->assign(v, v + n);
  m->beta_dirty = true;
  ECO_CATCH
}
int eco_get_f64(eco_model* m, const char* name, double* out, size_t cap, size_t* n) {
  ECO_TRY
  need(m);
  auto* a = f64_member(m, name);
  if (!a) throw std::runtime_error(std::string("unknown f64 member: ") + name);
  copy_out(*a, out, cap, n);
  ECO_CATCH
}
int eco_set_u32(eco_model* m, const char* name, const uint32_t* v, size_t n) {
  ECO_TRY
  need(m);
  const std::string s(name);
  if (s == "interaction.row") m->row.assign(v, v + n);
  else if (s == "interaction.col") m->col.assign(v, v + n);
  else throw std::runtime_error("unknown u32 member: " + s);
  m->beta_dirty = true;
  ECO_CATCH
}
int eco_get_u32(eco_model* m, const char* name, uint32_t* out, size_t cap, size_t* n) {
  ECO_TRY
  need(m);
  const std::string s(name);
  if (s == "interaction.row") copy_out(m->row, out, cap, n);
  else if (s == "interaction.col") copy_out(m->col, out, cap, n);
  else throw std::runtime_error("unknown u32 member: " + s);
  ECO_CATCH
}
int eco_set_i32(eco_model* m, const char* name, const int32_t* v, size_t n) {
  ECO_TRY
  need(m);
  if (std::string(name) != "species") throw std::runtime_error(std::string("unknown i32 member: ") + name);
  m->species.assign(v, v + n);
  m->beta_dirty = true;
  ECO_CATCH
}
int eco_get_i32(eco_model* m, const char* name, int32_t* out, size_t cap, size_t* n) {
  ECO_TRY
  need(m);
  if (std::string(name) != "species") throw std::runtime_error(std::string("unknown i32 member: ") + name);
  copy_out(m->species, out, cap, n);
  ECO_CATCH
}
int eco_set_scalar(eco_model* m, const char* name, double v) {
  ECO_TRY
  need(m);
  const std::string s(name);
  if (double* p = scalar_member(m, s)) *p = v;
  else if (s == "tstep") m->tstep = (uint64_t)v;
  else if (s == "last_mut_tstep") m->last_mut_tstep = (uint64_t)v;
  else if (s == "last_mig_tstep") m->last_mig_tstep = (uint64_t)v;
  else if (s == "fixMutation") m->fixMutation = v != 0;
  else if (s == "fixMigration") m->fixMigration = v != 0;
  else if (s == "nprocs") m->nprocs = (unsigned)v;
  else throw std::runtime_error("unknown scalar member: " + s);
  ECO_CATCH
}
int eco_get_scalar(eco_model* m, const char* name, double* out) {
  ECO_TRY
  need(m);
  const std::string s(name);
  if (double* p = scalar_member(m, s)) *out = *p;
  else if (s == "tstep") *out = (double)m->tstep;
  else if (s == "last_mut_tstep") *out = (double)m->last_mut_tstep;
  else if (s == "last_mig_tstep") *out = (double)m->last_mig_tstep;
  else if (s == "fixMutation") *out = m->fixMutation;
  else if (s == "fixMigration") *out = m->fixMigration;
  else if (s == "nsp") *out = (double)std::max(m->nsp(), (size_t)m->eng.nsp());
  else if (s == "nnz") *out = (double)m->val.size();
  else if (s == "ncells") *out = (double)(m->numX * m->numY);
  else if (s == "ncells_local") *out = (double)m->ncells_local();
  else if (s == "numX") *out = (double)m->numX;
  else if (s == "numY") *out = (double)m->numY;
  else if (s == "maxNbrs") *out = (double)m->maxNbrs;
  else if (s == "launches") *out = (double)m->eng.launches();
  else if (s == "bounds_bad") *out = (double)m->eng.bounds_bad();                      // bounds-assert build only (0 otherwise)
  else if (s == "bounds_selftest") *out = (double)m->eng.bounds_selftest();
  else if (s == "bounds_site") { unsigned st = 0; m->eng.bounds_bad(&st); *out = (double)st; }
  else if (s == "last_block") { int g, b, sm; m->eng.last_kernel_config(&g, &b, &sm); *out = (double)b; }   // threads per CTA of the last hot kernel
  else if (s == "last_smem") { int g, b, sm; m->eng.last_kernel_config(&g, &b, &sm); *out = (double)sm; }
  else if (s == "device_bytes") *out = (double)m->eng.bytes_allocated();
  else if (s == "dense_active") *out = m->eng.dense_wanted() ? 1.0 : 0.0;
  else if (s == "dense_q") *out = (double)m->eng.dense_last_q();
  else if (s == "dense_macs") *out = (double)m->eng.dense_mma_per_step();
  else if (s.rfind("dense_prof", 0) == 0) {   // dense_prof0..7: in-kernel cycle counters of CTA 0 (first call enables them)
    unsigned long long v[8];
    m->eng.dense_profile(v, true);
    *out = (double)v[std::min(7, std::max(0, s.size() > 10 ? s[10] - '0' : 0))];
  }
  else if (s == "dense_clock_mhz") {   // SM clock inside the last dense launch: cycles / ns of CTA 0's issue loop (first call enables it)
    unsigned long long v[8];
    m->eng.dense_profile(v, false);
    *out = v[7] ? (double)(v[0] + v[1] + v[2]) / (double)v[7] * 1e3 : 0.0;
  }
  else if (s == "nprocs") *out = (double)m->nprocs;
  else if (s == "y0") *out = (double)m->y0;
  else if (s == "ny_local") *out = (double)m->nyl;
  else if (s == "panmictic") *out = m->panmictic ? 1.0 : 0.0;
  else if (s == "buf_version") *out = (double)m->eng.buf_version();
  else if (s == "tile_width") *out = (double)m->eng.tile_width();     // cells per tile of the blocked layout (tests, tools)
  else if (s == "capacity") *out = (double)m->eng.capacity();
  else if (s == "peer_connected") *out = m->eng.peer_connected() ? 1.0 : 0.0;
  else throw std::runtime_error("unknown scalar member: " + s);
  ECO_CATCH
}

int eco_set_grid(eco_model* m, size_t nx, size_t ny) {
  ECO_TRY
  need(m);
  m->setGrid(nx, ny, 0, ny);
  ECO_CATCH
}
int eco_set_grid_slab(eco_model* m, size_t nx, size_t ny, size_t y0, size_t nyl) {
  ECO_TRY
  need(m);
  m->setGrid(nx, ny, y0, nyl);
  ECO_CATCH
}

int eco_set_density(eco_model* m, size_t cell, const int32_t* d, size_t n) {
  ECO_TRY
  need(m);
  const size_t lc = m->local_cell(cell);
  for (size_t i = 0; i < n; ++i)
    if (d[i] < 0) throw std::runtime_error("density must be non-negative (reference asserts density>=0, models/ecolab_model.cc:108)");
  // the panmictic model IS its one cell: assigning its density replaces the array, shorter or longer (pred-prey.py
  // re-initialises the global object with 2 species after any earlier run); a spatial cell is padded to the model's count
  if (m->panmictic && std::max(n, m->nsp()) < m->eng.nsp()) m->eng.resize_species((uint32_t)std::max(n, m->nsp()));
  m->grow_engine(std::max(n, m->nsp()));
  m->eng.set_density(lc, d, n);
  ECO_CATCH
}
int eco_get_density(eco_model* m, size_t cell, int32_t* out, size_t cap, size_t* n) {
  ECO_TRY
  need(m);
  const size_t lc = m->local_cell(cell);
  const size_t ns = m->eng.nsp();
  if (n) *n = ns;
  if (out && ns) {
    if (cap >= ns) m->eng.get_density(lc, out, ns);
    else {
      std::vector<int32_t> tmp(ns);
      m->eng.get_density(lc, tmp.data(), ns);
      std::copy(tmp.begin(), tmp.begin() + cap, out);
    }
  }
  ECO_CATCH
}
int eco_set_density_all(eco_model* m, const int32_t* d, size_t ncells, size_t nsp) {
  ECO_TRY
  need(m);
  if (ncells != m->ncells_local()) throw std::runtime_error("set_density_all: cell count differs from the owned grid");
  m->grow_engine(std::max(nsp, m->nsp()));
  if (nsp != m->eng.nsp()) throw std::runtime_error("set_density_all: species count differs from the model");
  for (size_t i = 0; i < ncells * nsp; ++i)
    if (d[i] < 0) throw std::runtime_error("density must be non-negative (reference asserts density>=0, models/ecolab_model.cc:108)");
  m->eng.set_density_all(d, nsp);
  ECO_CATCH
}
int eco_get_density_all(eco_model* m, int32_t* out, size_t ncells, size_t nsp) {
  ECO_TRY
  need(m);
  if (ncells != m->ncells_local()) throw std::runtime_error("get_density_all: cell count differs from the owned grid");
  if (nsp != m->eng.nsp()) throw std::runtime_error("get_density_all: species count differs from the model");
  m->eng.get_density_all(out, nsp);
  ECO_CATCH
}
int eco_get_density_rows(eco_model* m, size_t y, size_t nrows, int32_t* out, size_t nsp) {
  ECO_TRY
  need(m);
  if (y < m->y0 || y + nrows > m->y0 + m->nyl) throw std::runtime_error("get_density_rows: rows are not owned by this process");
  if (nsp != m->eng.nsp()) throw std::runtime_error("get_density_rows: species count differs from the model");
  m->eng.get_density_rows((uint32_t)(y - m->y0), (uint32_t)nrows, out, nsp);
  ECO_CATCH
}
int eco_set_density_rows(eco_model* m, size_t y, size_t nrows, const int32_t* d, size_t nsp) {
  ECO_TRY
  need(m);
  if (y < m->y0 || y + nrows > m->y0 + m->nyl) throw std::runtime_error("set_density_rows: rows are not owned by this process");
  m->grow_engine(std::max(nsp, m->nsp()));
  if (nsp != m->eng.nsp()) throw std::runtime_error("set_density_rows: species count differs from the model");
  for (size_t i = 0; i < nrows * m->numX * nsp; ++i)
    if (d[i] < 0) throw std::runtime_error("density must be non-negative (reference asserts density>=0, models/ecolab_model.cc:108)");
  m->eng.set_density_rows((uint32_t)(y - m->y0), (uint32_t)nrows, d, nsp);
  ECO_CATCH
}
int eco_set_density_all_device(eco_model* m, const int32_t* d, size_t ncells, size_t nsp) {
  ECO_TRY
  need(m);
  if (ncells != m->ncells_local()) throw std::runtime_error("set_density_all_device: cell count differs from the owned grid");
  m->grow_engine(std::max(nsp, m->nsp()));
  if (nsp != m->eng.nsp()) throw std::runtime_error("set_density_all_device: species count differs from the model");
  m->eng.set_density_all_dev(d, nsp);
  ECO_CATCH
}
int eco_get_density_all_device(eco_model* m, int32_t* out, size_t ncells, size_t nsp) {
  ECO_TRY
  need(m);
  if (ncells != m->ncells_local()) throw std::runtime_error("get_density_all_device: cell count differs from the owned grid");
  if (nsp != m->eng.nsp()) throw std::runtime_error("get_density_all_device: species count differs from the model");
  m->eng.get_density_all_dev(out, nsp);
  ECO_CATCH
}
int eco_fill_density(eco_model* m, int32_t v) {
  ECO_TRY
  need(m);
  if (v < 0) throw std::runtime_error("density must be non-negative");
  m->grow_engine(m->nsp());
  m->eng.fill_density(v);
  ECO_CATCH
}
int eco_fill_density_random(eco_model* m, int32_t lo, int32_t hi) {
  ECO_TRY
  need(m);
  if (lo < 0 || hi < lo) throw std::runtime_error("fill_density_random: need 0 <= lo <= hi");
  m->grow_engine(m->nsp());
  m->eng.fill_density_random(m->seed, lo, hi);
  ECO_CATCH
}
int eco_get_salt(eco_model* m, size_t cell, double* out) {
  ECO_TRY
  need(m);
  *out = (double)philox_host(m->seed, 2, (uint32_t)(cell % m->numX), (uint32_t)(cell / m->numX), 0, m->salt_t) / 4294967295.0;   // :619-620
  ECO_CATCH
}

int eco_make_consistent(eco_model* m) {
  ECO_TRY
  need(m);
  m->makeConsistent();
  ECO_CATCH
}
int eco_seed(eco_model* m, uint32_t s) {
  ECO_TRY
  need(m);
  m->seed = s;
  ECO_CATCH
}
int eco_generate(eco_model* m, unsigned niter) {
  ECO_TRY
  need(m);
  m->generate(niter);
  ECO_CATCH
}
int eco_migrate(eco_model* m, uint64_t* total) {
  ECO_TRY
  need(m);
  const uint64_t t = m->migrate_local();
  if (total) *total = t / 2;   // :702
  ECO_CATCH
}
int eco_step(eco_model* m, unsigned niter, uint64_t* total) {
  ECO_TRY
  need(m);
  m->generate(niter);
  const uint64_t t = m->migrate_local();
  if (total) *total = t / 2;
  ECO_CATCH
}
int eco_step_host(eco_model* m, const int32_t* in, int32_t* out, size_t ncells, size_t nsp, unsigned niter, uint64_t* total) {
  ECO_TRY
  need(m);
  if (!in || !out) throw std::runtime_error("step_host: null buffer");
  const uint64_t t = m->step_host(in, out, ncells, nsp, niter);
  if (total) *total = m->partitioned() ? t : t / 2;   // :702; a slab returns its unhalved share (summed over ranks by the caller)
  ECO_CATCH
}
int eco_step_async(eco_model* m, unsigned niter) {
  ECO_TRY
  need(m);
  m->generate(niter);
  m->migrate_local(false);
  ECO_CATCH
}
int eco_migrate_wait(eco_model* m, uint64_t* total) {
  ECO_TRY
  need(m);
  const uint64_t t = m->eng.migrate_result();
  if (total) *total = t;
  ECO_CATCH
}
int eco_step_host_begin(eco_model* m, const int32_t* in, size_t ncells, size_t nsp, unsigned niter) {
  ECO_TRY
  need(m);
  if (!in) throw std::runtime_error("step_host_begin: null buffer");
  m->step_host_begin(in, ncells, nsp, niter);
  ECO_CATCH
}
int eco_migrate_local(eco_model* m, uint64_t* total) {
  ECO_TRY
  need(m);
  const uint64_t t = m->migrate_local();
  if (total) *total = t;
  ECO_CATCH
}
int eco_migrate_local_async(eco_model* m, void** total_dev) {
  ECO_TRY
  need(m);
  if (!m->nsp()) throw std::runtime_error("migrate_local_async: no species");
  m->migrate_local(false);
  if (total_dev) *total_dev = m->eng.total_dev();
  ECO_CATCH
}
int eco_condense(eco_model* m, unsigned* removed) {
  ECO_TRY
  need(m);
  const unsigned r = m->condense();
  if (removed) *removed = r;
  ECO_CATCH
}
int eco_mutate(eco_model* m) {
  ECO_TRY
  need(m);
  m->mutate();
  ECO_CATCH
}
int eco_random_interaction(eco_model* m, unsigned conn, double sigma) {
  ECO_TRY
  need(m);
  m->random_interaction(conn, sigma);
  ECO_CATCH
}
int eco_connectivity(eco_model* m, double* out) {
  ECO_TRY
  need(m);
  *out = m->connectivity();
  ECO_CATCH
}
int eco_nsp_per_cell(eco_model* m, uint32_t* out, size_t ncells) {
  ECO_TRY
  need(m);
  if (ncells != m->ncells_local()) throw std::runtime_error("nsp_per_cell: cell count differs from the owned grid");
  m->eng.nsp_per_cell(out);
  ECO_CATCH
}
int eco_lifetimes(eco_model* m, double* out, size_t cap, size_t* n) {
  ECO_TRY
  need(m);
  if (!m->panmictic) throw std::runtime_error("lifetimes is a panmictic_ecolab method");
  const size_t ns = m->nsp();
  std::vector<int32_t> d(ns);
  if (ns) m->eng.get_density(0, d.data(), std::min<size_t>(ns, m->eng.nsp()));
  if (m->create.size() != ns) throw std::runtime_error("inconsistent model: create size differs from species");
  std::vector<double> lt;
  for (size_t i = 0; i < ns; ++i) {   // :478-488
    if (m->create[i] == 0 && d[i] > 10) m->create[i] = (double)m->tstep;
    else if (m->create[i] > 0 && d[i] == 0) { lt.push_back(m->tstep - m->create[i]); m->create[i] = 0; }
  }
  copy_out(lt, out, cap, n);
  ECO_CATCH
}
int eco_mutant_connectivity(eco_model* m, size_t before, double* out) {
  ECO_TRY
  need(m);
  const size_t ns = m->nsp();
  *out = 0;
  if (ns > before) {   // :496-501
    double s = 0;
    for (size_t j = 0; j < m->val.size(); ++j)
      if (m->row[j] > before || m->col[j] > before) s += (m->val[j] * m->val[j]) / (m->diag[m->row[j]] * m->diag[m->col[j]]);
    *out = s / ((ns - before) * (ns - 1));
  }
  ECO_CATCH
}
int eco_extinction_connectivity(eco_model* m, double* out) {
  ECO_TRY
  need(m);
  const size_t ns = m->nsp();
  std::vector<int32_t> d(ns);
  if (ns) m->eng.get_density(0, d.data(), std::min<size_t>(ns, m->eng.nsp()));
  size_t dead = 0;
  for (int v : d) dead += v == 0;
  *out = 0;
  if (dead) {   // :509-514
    double s = 0;
    for (size_t j = 0; j < m->val.size(); ++j)
      if (d[m->row[j]] == 0 || d[m->col[j]] == 0) s += (m->val[j] * m->val[j]) / (m->diag[m->row[j]] * m->diag[m->col[j]]);
    *out = s / (dead * (ns - 1));
  }
  ECO_CATCH
}
int eco_species_totals(eco_model* m, int64_t* out, size_t nsp) {
  ECO_TRY
  need(m);
  if (nsp != m->eng.nsp()) throw std::runtime_error("species_totals: species count differs from the model");
  m->eng.species_totals(out);
  ECO_CATCH
}
int eco_species_totals_noalloc(eco_model* m, int64_t* out, size_t nsp) {
  ECO_TRY
  need(m);
  if (nsp > eco::Engine::noalloc_max_species())
    throw std::runtime_error("species_totals_noalloc: " + std::to_string(nsp) + " species exceed the static scratch of " +
                             std::to_string(eco::Engine::noalloc_max_species()) + " (ECO_NOALLOC_MAX_SPECIES); use eco_species_totals");
  if (nsp != m->eng.nsp()) throw std::runtime_error("species_totals_noalloc: species count differs from the model");
  if (nsp && !out) throw std::runtime_error("species_totals_noalloc: null output");
  m->eng.species_totals_noalloc(out);
  ECO_CATCH
}

}  // extern "C"

// ---- checkpoint / restart: self-describing little-endian dump -------------------------
namespace {
const char kMagic[8] = {'E', 'C', 'O', 'B', '2', '0', '0', 3};
template <class T> void wr(FILE* f, const T& v) { if (fwrite(&v, sizeof v, 1, f) != 1) throw std::runtime_error("checkpoint: write failed"); }
template <class T> void rd(FILE* f, T& v) { if (fread(&v, sizeof v, 1, f) != 1) throw std::runtime_error("restart: truncated file"); }
template <class T> void wrv(FILE* f, const std::vector<T>& v) {
  wr(f, (uint64_t)v.size());
  if (!v.empty() && fwrite(v.data(), sizeof(T), v.size(), f) != v.size()) throw std::runtime_error("checkpoint: write failed");
}
template <class T> void rdv(FILE* f, std::vector<T>& v) {
  uint64_t n; rd(f, n);
  v.resize(n);
  if (n && fread(v.data(), sizeof(T), n, f) != n) throw std::runtime_error("restart: truncated file");
}
}  // namespace

extern "C" {

int eco_checkpoint(eco_model* m, const char* file) {
  ECO_TRY
  need(m);
  FILE* f = fopen(file, "wb");
  if (!f) throw std::runtime_error(std::string("checkpoint: cannot open ") + file);
  struct Closer { FILE* f; ~Closer() { fclose(f); } } closer{f};
  fwrite(kMagic, 1, 8, f);
  wr(f, (uint8_t)m->panmictic); wr(f, (uint64_t)m->numX); wr(f, (uint64_t)m->numY); wr(f, (uint64_t)m->y0); wr(f, (uint64_t)m->nyl);
  wr(f, (uint64_t)m->maxNbrs);
  wr(f, m->tstep); wr(f, m->last_mut_tstep); wr(f, m->last_mig_tstep);
  wr(f, m->sp_sep); wr(f, m->mut_max); wr(f, m->repro_min); wr(f, m->repro_max); wr(f, m->odiag_min); wr(f, m->odiag_max); wr(f, m->gen_bias);
  wr(f, (uint8_t)m->fixMutation); wr(f, (uint8_t)m->fixMigration);
  wr(f, m->seed); wr(f, m->salt_t); wr(f, (int32_t)m->sp_cntr);
  // the process-wide generators (array_urand, array_grand, src/arrays.cc:24-25) and the rank count that spaces the species
  // labels (:462): with them a restarted run also draws the same new-species parameters (the reference saves neither)
  wr(f, (uint32_t)m->nprocs); wr(f, g_grand.sum); wr(f, (uint32_t)g_grand.n);
  for (int32_t v : g_urand.r) wr(f, v);
  wr(f, g_urand.f); wr(f, g_urand.b);
  wrv(f, m->species); wrv(f, m->create); wrv(f, m->repro_rate); wrv(f, m->mutation); wrv(f, m->migration);
  wrv(f, m->diag); wrv(f, m->val); wrv(f, m->row); wrv(f, m->col);
  const size_t ns = m->eng.nsp();
  std::vector<int32_t> d(m->ncells_local() * ns);
  if (!d.empty()) m->eng.get_density_all(d.data(), ns);
  wr(f, (uint64_t)ns);
  wrv(f, d);
  ECO_CATCH
}
int eco_restart(eco_model* m, const char* file) {
  ECO_TRY
  need(m);
  FILE* f = fopen(file, "rb");
  if (!f) throw std::runtime_error(std::string("restart: cannot open ") + file);
  struct Closer { FILE* f; ~Closer() { fclose(f); } } closer{f};
  char magic[8];
  if (fread(magic, 1, 8, f) != 8 || memcmp(magic, kMagic, 8)) throw std::runtime_error("restart: not an ecolab_b200 checkpoint");
  uint8_t pan, fm, fg; uint64_t nx, ny, y0, nyl, mn;
  rd(f, pan); rd(f, nx); rd(f, ny); rd(f, y0); rd(f, nyl); rd(f, mn);
  rd(f, m->tstep); rd(f, m->last_mut_tstep); rd(f, m->last_mig_tstep);
  rd(f, m->sp_sep); rd(f, m->mut_max); rd(f, m->repro_min); rd(f, m->repro_max); rd(f, m->odiag_min); rd(f, m->odiag_max); rd(f, m->gen_bias);
  rd(f, fm); rd(f, fg);
  m->fixMutation = fm; m->fixMigration = fg;
  int32_t spc;
  rd(f, m->seed); rd(f, m->salt_t); rd(f, spc);
  m->sp_cntr = spc;
  uint32_t np, gn;
  rd(f, np); rd(f, g_grand.sum); rd(f, gn);
  m->nprocs = np; g_grand.n = gn;
  for (int32_t& v : g_urand.r) rd(f, v);
  rd(f, g_urand.f); rd(f, g_urand.b);
  if (g_urand.f < 0 || g_urand.f > 30 || g_urand.b < 0 || g_urand.b > 30) throw std::runtime_error("restart: corrupt generator state");
  rdv(f, m->species); rdv(f, m->create); rdv(f, m->repro_rate); rdv(f, m->mutation); rdv(f, m->migration);
  rdv(f, m->diag); rdv(f, m->val); rdv(f, m->row); rdv(f, m->col);
  uint64_t ns; rd(f, ns);
  std::vector<int32_t> d; rdv(f, d);
  for (int32_t v : d)
    if (v < 0) throw std::runtime_error("restart: negative density in the checkpoint");
  if (pan) {
    m->panmictic = true; m->numX = m->numY = 1; m->y0 = 0; m->nyl = 1;
    m->eng.set_grid(1, 1, 0, 1, (uint32_t)ns);
  } else {
    m->setGrid(nx, ny, y0, nyl);
    m->eng.resize_species((uint32_t)ns);
  }
  m->maxNbrs = mn;
  if (ns > m->eng.nsp()) m->eng.resize_species((uint32_t)ns);
  if (!d.empty()) m->eng.set_density_all(d.data(), ns);
  m->beta_dirty = true;
  ECO_CATCH
}

// ---- multi-GPU halves ---------------------------------------------------------------------
int eco_halo_ptrs(eco_model* m, void** st, void** sb, void** rt, void** rb, size_t* bytes) {
  ECO_TRY
  need(m);
  m->eng.halo_ptrs(st, sb, rt, rb);
  if (bytes) *bytes = m->eng.row_bytes();
  ECO_CATCH
}
int eco_peer_export(eco_model* m, void* blob, size_t blob_bytes) {
  ECO_TRY
  need(m);
  static_assert(sizeof(eco::Engine::PeerBlob) <= ECO_PEER_BLOB_BYTES, "peer blob size");
  if (blob_bytes < ECO_PEER_BLOB_BYTES) throw std::runtime_error("peer_export: blob buffer too small");
  memset(blob, 0, ECO_PEER_BLOB_BYTES);
  m->eng.peer_export(reinterpret_cast<eco::Engine::PeerBlob*>(blob));
  ECO_CATCH
}
int eco_peer_connect(eco_model* m, const void* top_blob, const void* bottom_blob) {
  ECO_TRY
  need(m);
  m->eng.peer_connect(reinterpret_cast<const eco::Engine::PeerBlob*>(top_blob), reinterpret_cast<const eco::Engine::PeerBlob*>(bottom_blob));
  ECO_CATCH
}
int eco_peer_disconnect(eco_model* m) {
  ECO_TRY
  need(m);
  m->eng.peer_disconnect();
  ECO_CATCH
}
int eco_alive_mask(eco_model* m, uint8_t* alive, size_t nsp) {
  ECO_TRY
  need(m);
  if (nsp != m->nsp() || nsp != m->eng.nsp()) throw std::runtime_error("alive_mask: species count differs from the model");
  if (nsp) m->eng.alive_mask(alive);
  ECO_CATCH
}
int eco_alive_mask_device(eco_model* m, void** mask_dev, size_t nsp) {
  ECO_TRY
  need(m);
  if (!mask_dev) throw std::runtime_error("alive_mask_device: null output");
  if (nsp != m->nsp() || nsp != m->eng.nsp()) throw std::runtime_error("alive_mask_device: species count differs from the model");
  *mask_dev = nsp ? m->eng.alive_mask_device() : nullptr;
  ECO_CATCH
}
int eco_condense_apply(eco_model* m, const uint8_t* alive, size_t nsp, unsigned* removed) {
  ECO_TRY
  need(m);
  if (nsp != m->nsp() || nsp != m->eng.nsp()) throw std::runtime_error("condense_apply: species count differs from the model");
  const unsigned r = m->condense_apply(std::vector<uint8_t>(alive, alive + nsp));
  if (removed) *removed = r;
  ECO_CATCH
}
int eco_mutate_cells(eco_model* m, size_t* n) {
  ECO_TRY
  need(m);
  m->mutate_cells(m->pending);
  if (n) *n = m->pending.size();
  ECO_CATCH
}
int eco_mutate_records(eco_model* m, eco_mut_record* out, size_t cap) {
  ECO_TRY
  need(m);
  if (cap < m->pending.size()) throw std::runtime_error("mutate_records: output buffer too small");
  static_assert(sizeof(eco_mut_record) == sizeof(eco::MutRecord), "record layout");
  if (!m->pending.empty()) memcpy(out, m->pending.data(), m->pending.size() * sizeof(eco_mut_record));
  ECO_CATCH
}
int eco_mutate_model(eco_model* m, const uint32_t* parents, size_t n) {
  ECO_TRY
  need(m);
  m->check_sizes(false);
  if (m->mutation.size() != m->nsp() || m->migration.size() != m->nsp() || m->create.size() != m->nsp())
    throw std::runtime_error("inconsistent model: mutation/migration/create size differs from species");
  m->mutateModel(std::vector<uint32_t>(parents, parents + n));
  ECO_CATCH
}
int eco_append_species(eco_model* m, const uint32_t* cells, size_t n) {
  ECO_TRY
  need(m);
  m->eng.append_species(cells, n);
  ECO_CATCH
}

void eco_array_urand_seed(int seed) { g_urand.seed((unsigned)seed); }
void eco_array_urand_rand(double* out, size_t n) { for (size_t k = 0; k < n; ++k) out[k] = urand01(); }
void eco_array_grand_reset(void) { g_grand = RunningClt(); }

int eco_set_tuning(eco_model* m, const char* knob, int value) {
  ECO_TRY
  need(m);
  const std::string k(knob);
  if (k == "timing") { m->eng.reset_kernel_times(); m->eng.timing = value != 0; }
  else if (k == "mig_rows") m->eng.mig_rows = value;
  else if (k == "gen_stages") m->eng.gen_stages = value;
  else if (k == "gen_nobsmem") m->eng.gen_nobsmem = value;
  else if (k == "gen_dual") m->eng.gen_dual = value;
  else if (k == "gen_dense") m->eng.gen_dense = value;
  else if (k == "dense_slices") {
    if (value < 1 || value > 7) throw std::runtime_error("dense_slices must be 1..7");
    m->eng.dense_slices = value;
  }
  else throw std::runtime_error("unknown tuning knob: " + k);
  ECO_CATCH
}

int eco_kernel_times(eco_model* m, double* out, size_t cap, size_t* n) {
  ECO_TRY
  need(m);
  double ms[eco::Engine::K_COUNT];
  uint64_t cnt[eco::Engine::K_COUNT];
  m->eng.kernel_times(ms, cnt);
  std::vector<double> v;
  for (int k = 0; k < eco::Engine::K_COUNT; ++k) { v.push_back(ms[k]); v.push_back((double)cnt[k]); }
  copy_out(v, out, cap, n);
  ECO_CATCH
}
int eco_pre_rounding(eco_model* m, double* out, size_t ncells, size_t nsp) {
  ECO_TRY
  need(m);
  if (m->tstep == 0) m->makeConsistent();
  m->check_sizes(false);
  if (ncells != m->ncells_local() || nsp != m->nsp()) throw std::runtime_error("pre_rounding: shape differs from the model");
  m->upload();
  m->eng.pre_rounding(out);
  ECO_CATCH
}

}  // extern "C"
