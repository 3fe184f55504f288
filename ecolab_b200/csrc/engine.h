// engine.h -- internal interface between the host model (model.cpp, the C ABI)
// and the device store + sm_100a kernels (engine.cu).  Not a public header; the
// public boundary is include/ecolab_b200.h.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace eco {

// One mutation event produced by the per-cell speciation kernel:
// `count` mutants of parent species slot `slot` were born in cell `cell`.
struct MutRecord {
  uint32_t cell, slot, count;
};

// beta in CSR form, rows in species order, the entries of a row in COO order (reference oDiagIdx,
// models/ecolab_model.cc:718-723) so that the FP64 summation order is the reference's.  The engine
// re-blocks it into the ELL form its kernels read (k_generate.cuh).
struct BetaCsr {
  uint32_t nsp = 0;
  std::vector<uint32_t> rowptr;   // [nsp+1]
  std::vector<double> val;        // [nnz]
  std::vector<uint32_t> col;      // [nnz]
  std::vector<double> diag, rr;   // [nsp] interaction.diag, repro_rate
};

class Engine {
 public:
  explicit Engine(int device);
  ~Engine();

  // ---- geometry / storage -------------------------------------------------
  // This process owns grid rows [y0, y0+ny) of an nx x ny_global grid
  // (single GPU: y0=0, ny=ny_global).  Density is reset to zero.
  void set_grid(uint32_t nx, uint32_t ny_global, uint32_t y0, uint32_t ny, uint32_t nsp);
  // Change the species count (keeps existing rows; new rows are zero).
  void resize_species(uint32_t nsp);
  uint32_t nsp() const { return nsp_; }
  uint32_t nx() const { return nx_; }
  uint32_t ny() const { return ny_; }
  uint32_t y0() const { return y0_; }
  uint32_t ny_global() const { return nyg_; }
  size_t ncells_local() const { return (size_t)nx_ * ny_; }

  void set_stream(void* s);
  void* stream() const { return (void*)stream_; }
  void* total_dev() const { return (void*)d_total_; }   // device copy of the last migrate's sum |delta| (uint64)
  void sync();

  // density access, cells addressed by LOCAL index x + nx*(y-y0)
  void set_density(size_t cell, const int32_t* d, size_t n);
  void get_density(size_t cell, int32_t* d, size_t n);
  // all local cells, host buffer cell-major [cell][nsp]; goes through a pinned/
  // device staging buffer and a transposing kernel
  void set_density_all(const int32_t* host, size_t nsp);
  void get_density_all(int32_t* host, size_t nsp);
  // owned grid rows [r0, r0+nrows) (LOCAL row index), host buffer cell-major [cell][nsp]
  void get_density_rows(uint32_t r0, uint32_t nrows, int32_t* host, size_t nsp);
  void set_density_rows(uint32_t r0, uint32_t nrows, const int32_t* host, size_t nsp);
  // same with a DEVICE cell-major buffer (used by bench to build big states)
  void set_density_all_dev(const int32_t* dev, size_t nsp);
  void get_density_all_dev(int32_t* dev, size_t nsp);
  void fill_density(int32_t v);
  void fill_density_random(uint32_t seed, int32_t lo, int32_t hi);   // lo + philox(stage 3) % (hi-lo+1)

  // ---- species data ---------------------------------------------------------
  void upload_beta(BetaCsr b);

  // ---- hot path -------------------------------------------------------------
  // niter fused Lotka-Volterra steps on every owned cell; t0 = absolute tstep of the first
  void generate(uint32_t niter, uint64_t t0, uint32_t seed);
  void generate_rows(uint32_t niter, uint64_t t0, uint32_t seed, uint32_t r0, uint32_t nrows);
  // cm[s] = capped migration coefficient; returns sum |delta| over owned cells (not halved)
  uint64_t migrate(const double* cm_host, bool wait_result = true);
  uint64_t migrate_result();   // blocks on the async total of the last migrate
  // generate(niter) + migrate with HOST cell-major buffers [cell][nsp] in and out, streamed in bands of grid rows:
  // H2D of band k+1, the kernels of band k and D2H of band k-1 overlap (full-duplex PCIe); returns sum |delta|
  uint64_t step_host(const int32_t* in_host, int32_t* out_host, uint32_t niter, uint64_t t0, uint32_t seed, const double* cm_host,
                     bool boundary_done = false);
  // partitioned grids: upload + generate the slab's first and last row ahead of the halo exchange (see engine_ops.inc)
  void step_host_begin(const int32_t* in_host, uint32_t niter, uint64_t t0, uint32_t seed);
  // alive[s] = 1 iff some owned cell has density != 0
  void alive_mask(uint8_t* alive_host);
  uint8_t* alive_mask_device();                        // the same, left on the device (stream-ordered)
  // compact species rows: keep[s] != 0 survives, order preserved
  void compact_species(const uint8_t* keep_host, uint32_t nkeep);
  // speciation: k = ROUND(ms[s] * n), n -= k; returns records sorted by (cell, slot); cell ids GLOBAL
  void mutate_cells(const double* ms_host, uint64_t t, uint32_t seed, std::vector<MutRecord>& out);
  // append one species per (global) cell id in `cells`: density 1 in that cell if owned, else 0
  void append_species(const uint32_t* cells, size_t m);
  // number of non-zero densities per owned cell
  void nsp_per_cell(uint32_t* out_host);
  // per-species totals over owned cells
  void species_totals(int64_t* out_host);
  // the same with the accumulators in a static device array: no cudaMalloc / cudaFree, at most noalloc_max_species() species;
  // `out` may be host or device memory.  Calls that overlap on one GPU race on the shared array.
  void species_totals_noalloc(int64_t* out);
  static uint32_t noalloc_max_species();

  // ---- multi-GPU halo -------------------------------------------------------
  // Device pointers of the boundary rows to send and the ghost rows to receive
  // (each row_bytes() long, contiguous).  Valid until the next migrate()/resize.
  void halo_ptrs(void** send_top, void** send_bottom, void** recv_top, void** recv_bottom);
  size_t row_bytes() const { return (size_t)ntx_ * cap_ * tw_ * sizeof(int32_t); }
  uint32_t tile_width() const { return tw_; }
  uint32_t capacity() const { return cap_; }
  // Halo over peer memory (CUDA IPC, NVLink): once connected, migrate's boundary units copy the neighbour slabs'
  // boundary rows straight out of the neighbours' buffers -- no ghost-row exchange, no collective.  Ordering is a
  // device-side epoch counter per rank: migrate() first publishes "everything before this migrate is done" (epoch e),
  // and a boundary unit waits until the neighbour's counter has reached e before it touches the neighbour's row.
  // The neighbour's counter reaching e also means its migrate e-1 has finished reading OUR rows, which is what makes
  // overwriting the destination buffer safe.  Every rank must call migrate the same number of times.
  struct PeerBlob {
    unsigned char handle[3][64];   // cudaIpcMemHandle_t of buffer 0, buffer 1, the flag words
    uint32_t ny, cap, tw, ntx;     // geometry of the exporting slab (its last owned row depends on ny)
    uint64_t version;
  };
  void peer_export(PeerBlob* out);                                  // allocates the second buffer if needed
  void peer_connect(const PeerBlob* top, const PeerBlob* bottom);   // neighbour above (lower y) / below; nullptr = none
  void peer_disconnect();
  bool peer_connected() const { return peer_open_[0] || peer_open_[1]; }
  uint64_t buf_version() const { return buf_version_; }            // bumped whenever a density buffer is (re)allocated

  // x = n + n*(r+ir) before ROUND for one generate sub-step, cell-major [cell][nsp]; densities untouched
  void pre_rounding(double* out_host);

  // per-kernel device timing (CUDA events on the engine stream), off by default
  enum KernelKind { K_GENERATE = 0, K_MIGRATE = 1, K_STEP = 2, K_DENSE_PACK = 3, K_DENSE_MMA = 4, K_COUNT = 5 };
  bool timing = false;
  void kernel_times(double* ms, uint64_t* count);   // arrays of K_COUNT; synchronises the stream
  void reset_kernel_times();

  // diagnostics
  uint64_t launches() const { return launches_; }
  void reset_launches() { launches_ = 0; }
  size_t bytes_allocated() const;
  int device() const { return device_; }
  void last_kernel_config(int* grid, int* block, int* smem) const {
    *grid = last_grid_; *block = last_block_; *smem = last_smem_;
  }
  // bounds-assert build (-DECO_BOUNDS): registers the ranges the kernels may touch / reads the violation count; no-ops otherwise
  void bounds_sync(const void* extra = nullptr, size_t extra_bytes = 0);
  unsigned bounds_bad(unsigned* site = nullptr, unsigned long long* addr = nullptr);
  unsigned bounds_selftest();
  // tuning knobs (bench/profiling): 0 = auto.  gen_stages: shared-memory tile buffers of the generate pipeline
  int gen_stages = 0;
  int gen_nobsmem = 0;   // 1: never stage beta in shared memory
  int gen_dual = 0;      // generate(1): two species-row blocks per consumer warp at a time; 0 / 1 = where applicable, -1 = off
  int mig_rows = 0;      // grid rows per migrate work unit; 0 = automatic
  // dense (tensor-core) generate, k_dense.cuh.  gen_dense: 1 = use it, 0 / -1 = never (default: it is not bit-exact against
  // the reference's FP64 summation order, so the caller must ask for it).  dense_slices = P, the 8-bit slices beta is cut into (3: 2^-24 of the row maximum; 7: FP64).
  int gen_dense = 0;
  int dense_slices = 3;
  bool dense_wanted() const;
  // CTA 0's counters of the last dense launch; the first call switches them on: [2] cycles and [7] ns of the issue loop always,
  // the per-stage split ([0] wait TMEM, [1] wait stage, [2] issue, [4..6] epilogue) only with detail (it slows the issuing thread)
  void dense_profile(unsigned long long* out8, bool detail);
  uint32_t dense_last_q() const { return dense_q_; }        // density slices used by the last dense generate
  uint64_t dense_mma_per_step() const { return dense_macs_; }   // integer MACs issued by the last dense generate step

 private:
  void ensure_other();
  void migrate_prepare(const double* cm_host);
  void migrate_rows(uint32_t r0, uint32_t nrows);
  void alloc_buffers(uint32_t cap);
  void build_ell();
  void build_dense();
  void generate_dense(uint32_t niter, uint64_t t0, uint32_t seed, uint32_t r0, uint32_t nrows, double* xout_dev);
  void ensure_vec(size_t n);
  void ensure_mask(size_t n);
  void free_buffers();
  int32_t* row0(int which) const;   // pointer to owned row 0 of buffer `which`
  size_t tile_stride() const { return (size_t)cap_ * tw_; }
  size_t row_stride() const { return (size_t)ntx_ * cap_ * tw_; }

  int device_ = 0;
  void* stream_ = nullptr;
  bool own_stream_ = false;
  int num_sms_ = 148;
  int max_smem_ = 0;

  uint32_t nx_ = 1, nyg_ = 1, y0_ = 0, ny_ = 1, ntx_ = 1;
  uint32_t nsp_ = 0, cap_ = 0, tw_ = 32;   // live species rows, allocated rows, tile width (cells)
  int32_t* buf_[2] = {nullptr, nullptr};   // allocations (with ghost rows)
  int cur_ = 0;
  // peer halo: [buffer index][side: 0 = above, 1 = below] boundary-row pointers inside the neighbours' buffers
  const void* peer_[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
  const unsigned int* peer_flag_[2] = {nullptr, nullptr};
  void* peer_map_[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};   // opened IPC mappings per side
  bool peer_open_[2] = {false, false};
  unsigned int* d_flags_ = nullptr;      // [0] epoch published by this rank, [2] migrate's peer-wait error word, [3] the dense kernel's watchdog word; IPC-exported
  unsigned int gen_epoch_ = 0;
  uint64_t buf_version_ = 0;

  // beta on device
  uint32_t nblocks_ = 0, beta_tw_ = 0, beta_nsp_ = 0;   // ELL blocks; the tile width / species count it was built for
  uint32_t* d_boff_ = nullptr; void* d_stream_ = nullptr;   // blocked-ELL stream (k_generate.cuh)
  size_t beta_cap_slots_ = 0, beta_cap_blocks_ = 0, beta_slots_ = 0;
  void* h_csr_ = nullptr; void* d_csr_ = nullptr; size_t csr_cap_ = 0;   // CSR staging (pinned) and its device copy, input of ell_build_kernel
  bool beta_fma_ok_ = true;    // every beta entry is zero or a normal number below 2^971: the exact-FMA product applies
  uint32_t beta_uslots_ = 0;   // != 0: every block of the stream has this many slots (uniform width, no offset table reads)
  BetaCsr beta_host_;                                   // kept to re-block when the tile width changes

  // dense path operands (k_dense.cuh)
  uint8_t* d_aop_ = nullptr; size_t aop_cap_ = 0;       // beta slices, shared-memory image layout
  uint8_t* d_bop_ = nullptr; size_t bop_cap_ = 0;       // density slices of one row band
  double* d_dsp_ = nullptr; size_t dsp_cap_ = 0;        // diag | repro_rate | 2^-s per padded species row
  double* d_partial_ = nullptr; size_t partial_cap_ = 0;
  unsigned int* d_maxn_ = nullptr; unsigned int* h_maxn_ = nullptr;
  uint32_t dense_p_ = 0, dense_nsp_ = 0, dense_mtiles_ = 0, dense_nkc_ = 0, dense_q_ = 1;
  uint64_t dense_macs_ = 0;
  unsigned long long* dense_prof_ = nullptr; bool dense_prof_detail_ = false;
  uint64_t beta_version_ = 0, dense_version_ = ~0ull;

  // scratch
  double* d_vec_ = nullptr; size_t vec_cap_ = 0;       // per-species doubles (cm / ms)
  uint8_t* d_mask_ = nullptr; uint32_t* d_map_ = nullptr; size_t mask_cap_ = 0;
  unsigned long long* d_total_ = nullptr;              // migrate total / counters
  unsigned long long* h_total_ = nullptr;              // pinned
  MutRecord* d_rec_ = nullptr; size_t rec_cap_ = 0;
  uint32_t* d_scan_ = nullptr; size_t scan_cap_ = 0;   // mutate: per-cell event counts + scan block totals
  int32_t* d_stage_ = nullptr; size_t stage_cap_ = 0;  // cell-major staging
  void* d_scratch_ = nullptr; size_t scratch_cap_ = 0;   // small per-call device temporaries (birth cells, species totals)
  void* scratch(size_t bytes);
  void relayout(uint32_t new_cap, uint32_t nsp);   // densities into a fresh buffer of that capacity (tile width follows)
  void retile_after_shrink();                      // condense: back to a wider tile when the survivors fit one
  int32_t* d_band_[4] = {nullptr, nullptr, nullptr, nullptr}; size_t band_cap_ = 0;   // step_host: 2 in + 2 out band stagings
  void* st_in_ = nullptr; void* st_out_ = nullptr;       // copy streams of step_host
  void* ev_band_[8] = {nullptr};
  void* ev_total_ = nullptr;

  uint64_t launches_ = 0;
  int last_grid_ = 0, last_block_ = 0, last_smem_ = 0;

  struct PendingEvent { void* start; void* stop; int kind; };
  std::vector<PendingEvent> ev_pending_;
  std::vector<void*> ev_pool_;
  double kt_ms_[K_COUNT] = {};
  uint64_t kt_n_[K_COUNT] = {};
  void time_begin(int kind);
  void time_end();
};

void check_cuda(int err, const char* what, const char* file, int line);

}  // namespace eco
