/* ecolab_b200.h -- C ABI of the B200-native EcoLab population-update engine.
 *
 * This is the drop-in boundary for ONE path of highperformancecoder/ecolab: the
 * ecology model's per-timestep update in models/ecolab_model.cc (generate /
 * mutate / migrate / condense) plus the model members the driver scripts read
 * and write.  The reference exposes that path as classdesc-reflected C++
 * objects (`panmictic_ecolab`, `spatial_ecolab`, models/ecolab_model.cc:32-41);
 * every entry point below names the reference member it replaces (file:line
 * relative to the reference checkout).  INTEGRATION.md shows the binding a
 * reference maintainer would add on top of this header.
 *
 * Conventions
 *   - plain C types only; host buffers are caller-owned, device memory is
 *     callee-owned; a model handle is bound to one CUDA device.
 *   - every function returns 0 on success, non-zero on failure;
 *     eco_last_error() returns the message of the calling thread's last failure
 *     (the reference throws C++ exceptions that surface as Python exceptions,
 *     include/error.h:24-49).
 *   - calls are synchronous at the API unless stated otherwise; one driver
 *     thread per handle (the reference is driven under the GIL).
 *   - the engine requires a CUDA device: eco_create fails loudly without one.
 *     There is no CPU fallback.
 *   - RNG: counter-based Philox4x32-10; for cell (x, y), species slot s, timestep t:
 *     block = philox(ctr={(x>>2) | stage<<30, y, s, t_lo}, key={seed, t_hi}), draw = block[x & 3];
 *     stage 0=generate 1=mutate 2=salt 3=synthetic fill (DESIGN.md "RNG contract").
 */
#ifndef ECOLAB_B200_H
#define ECOLAB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct eco_model eco_model;

/* record of one speciation event batch: `count` mutants of parent `slot` born in global cell `cell` */
typedef struct eco_mut_record { uint32_t cell, slot, count; } eco_mut_record;

const char* eco_last_error(void);
const char* eco_version(void);

/* ---- lifetime ------------------------------------------------------------------
 * A fresh model is panmictic (one cell; reference PanmicticModel,
 * models/ecolab_model.h:105-121) until eco_set_grid is called (SpatialModel,
 * models/ecolab_model.h:128-151). */
int eco_create(eco_model** out, int cuda_device);
int eco_destroy(eco_model* m);
/* run all kernels of this model on an existing CUDA stream (cudaStream_t) */
int eco_set_stream(eco_model* m, void* cuda_stream);
int eco_sync(eco_model* m);                          /* EcolabGraph::syncThreads, include/ecolab.h:240 */

/* ---- ModelData members (models/ecolab_model.h:47-57) ----------------------------
 * f64 arrays : "create" "repro_rate" "mutation" "migration"
 *              "interaction.diag" "interaction.val"       (include/sparse_mat.h:27)
 * u32 arrays : "interaction.row" "interaction.col"          (include/sparse_mat.h:29)
 * i32 array  : "species"
 * scalars    : "tstep" "last_mut_tstep" "last_mig_tstep" "sp_sep" "mut_max"
 *              "repro_min" "repro_max" "odiag_min" "odiag_max" "gen_bias"
 *              "fixMutation" "fixMigration"; read-only: "nsp" "nnz" "ncells"
 *              "numX" "numY" "maxNbrs" "launches" "device_bytes"
 * get_* with out==NULL returns the length in *n. */
int eco_set_f64(eco_model* m, const char* member, const double* v, size_t n);
int eco_get_f64(eco_model* m, const char* member, double* out, size_t cap, size_t* n);
int eco_set_u32(eco_model* m, const char* member, const uint32_t* v, size_t n);
int eco_get_u32(eco_model* m, const char* member, uint32_t* out, size_t cap, size_t* n);
int eco_set_i32(eco_model* m, const char* member, const int32_t* v, size_t n);
int eco_get_i32(eco_model* m, const char* member, int32_t* out, size_t cap, size_t* n);
int eco_set_scalar(eco_model* m, const char* member, double v);
int eco_get_scalar(eco_model* m, const char* member, double* out);

/* ---- grid (SpatialModel::setGrid, models/ecolab_model.cc:580-605) ----------------
 * cell id = x + nx*y (makeId, models/ecolab_model.h:137); von Neumann
 * neighbourhood, open boundaries; density reset to zero.  */
int eco_set_grid(eco_model* m, size_t nx, size_t ny);
/* multi-GPU: this process owns global rows [y0, y0+ny_local) of the nx x ny grid
 * (replaces graphcode's partition/distribute, graphcode/graphcode.h:475-482) */
int eco_set_grid_slab(eco_model* m, size_t nx, size_t ny, size_t y0, size_t ny_local);

/* ---- density (EcolabPoint::density, models/ecolab_model.h:88) ---------------------
 * `cell` is the GLOBAL cell id and must be owned by this process. */
int eco_set_density(eco_model* m, size_t cell, const int32_t* d, size_t n);
int eco_get_density(eco_model* m, size_t cell, int32_t* out, size_t cap, size_t* n);
/* all owned cells, host buffer cell-major [cell][nsp] in ascending cell id */
int eco_set_density_all(eco_model* m, const int32_t* d, size_t ncells, size_t nsp);
int eco_get_density_all(eco_model* m, int32_t* out, size_t ncells, size_t nsp);
/* a band of whole grid rows [y, y+nrows) (GLOBAL row index, all owned by this process), cell-major [cell][nsp] */
int eco_get_density_rows(eco_model* m, size_t y, size_t nrows, int32_t* out, size_t nsp);
int eco_set_density_rows(eco_model* m, size_t y, size_t nrows, const int32_t* d, size_t nsp);
/* same with DEVICE buffers (device-resident pipelines, benchmarks) */
int eco_set_density_all_device(eco_model* m, const int32_t* d_dev, size_t ncells, size_t nsp);
int eco_get_density_all_device(eco_model* m, int32_t* out_dev, size_t ncells, size_t nsp);
int eco_fill_density(eco_model* m, int32_t value);
/* synthetic state for benchmarks / full-size property tests (the scripts' `int(100*random())` initialisation,
 * models/spatial_ecolab.py:44): n[cell][slot] = lo + draw % (hi-lo+1), draw = Philox stage 3 at t=0, current seed */
int eco_fill_density_random(eco_model* m, int32_t lo, int32_t hi);
int eco_get_salt(eco_model* m, size_t cell, double* out);   /* EcolabPoint::salt, models/ecolab_model.h:87 */

/* ---- the hot path ----------------------------------------------------------------- */
/* ModelData::makeConsistent :705-716, SpatialModel::makeConsistent :725-735 */
int eco_make_consistent(eco_model* m);
/* PanmicticModel::seed / SpatialModel::seed, models/ecolab_model.h:111,144: sets the Philox seed key */
int eco_seed(eco_model* m, uint32_t seed);
/* SpatialModel::generate :607-614 / PanmicticModel::generate :49-54 (EcolabPoint::generate :90-112) */
int eco_generate(eco_model* m, unsigned niter);
/* SpatialModel::migrate :616-703; *total = sum|delta|/2 over the owned cells' deltas
 * (single GPU: the reference's return value; multi GPU: sum the UNHALVED locals, see eco_migrate_local) */
int eco_migrate(eco_model* m, uint64_t* total);
/* one driver-loop body of the spatial scripts (models/spatial_ecolab.py:67-69 minus mutate):
 * eco_generate(niter) followed by eco_migrate(total) as one call (fewer host round trips) */
int eco_step(eco_model* m, unsigned niter, uint64_t* total);
/* eco_step without the host round trip for migrate's return value: both kernels are enqueued on the model's stream and
 * the call returns; eco_migrate_wait blocks until the last migrate has finished and returns its sum|delta| (NOT halved,
 * owned cells only).  For throughput loops that read the total once at the end. */
int eco_step_async(eco_model* m, unsigned niter);
int eco_migrate_wait(eco_model* m, uint64_t* abs_delta_sum);
/* the same body for a HOST-resident state: densities come from `in` and the result goes to `out` (both cell-major
 * [cell][nsp], ascending cell id, as eco_set/get_density_all; pinned memory makes the copies asynchronous).  Equivalent
 * to eco_set_density_all(in); eco_step(niter); eco_get_density_all(out) -- i.e. SpatialModel::generate :607-614 +
 * ::migrate :616-703 on densities set through EcolabPoint::density (models/ecolab_model.h:88) -- but streamed in
 * bands of grid rows so that H2D, the kernels and D2H overlap.
 * On a slab of a partitioned grid (eco_set_grid_slab) the step has two halves around the halo exchange: eco_step_host_begin
 * uploads and generates the slab's first and last grid row, the host language then exchanges the halo rows (eco_halo_ptrs +
 * send/recv, or nothing once eco_peer_connect is live), and eco_step_host streams the rest; *total is then this slab's
 * UNHALVED sum|delta| (as eco_migrate_local). */
int eco_step_host_begin(eco_model* m, const int32_t* in, size_t ncells, size_t nsp, unsigned niter);
int eco_step_host(eco_model* m, const int32_t* in, int32_t* out, size_t ncells, size_t nsp, unsigned niter, uint64_t* total);
/* SpatialModel::condense :171-196 / PanmicticModel::condense :161-169; *removed = #extinct species */
int eco_condense(eco_model* m, unsigned* removed);
/* SpatialModel::mutate :215-276 / PanmicticModel::mutate :205-213 */
int eco_mutate(eco_model* m);
/* ModelData::random_interaction :43-47 (sparse_mat::init_rand, include/sparse_mat.h:88-111) */
int eco_random_interaction(eco_model* m, unsigned conn, double sigma);
/* ModelData::connectivity :317-323 */
int eco_connectivity(eco_model* m, double* out);
/* SpatialModel::nsp :117-122 (EcolabPoint::nsp :114-115): living species per owned cell */
int eco_nsp_per_cell(eco_model* m, uint32_t* out, size_t ncells);
/* PanmicticModel::lifetimes :474-490 */
int eco_lifetimes(eco_model* m, double* out, size_t cap, size_t* n);
/* PanmicticModel::mutantConnectivity :493-503, ::extinctionConnectivity :506-516 */
int eco_mutant_connectivity(eco_model* m, size_t before_nsp, double* out);
int eco_extinction_connectivity(eco_model* m, double* out);
/* per-species totals over owned cells (conservation checks, reference :644,:662) */
int eco_species_totals(eco_model* m, int64_t* out, size_t nsp);
/* No counterpart in the reference: added for callers that may not reach cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as eco_species_totals, but the
 * kernel's per-species accumulators live in a static device array of ECO_NOALLOC_MAX_SPECIES entries (128 KiB, held by
 * every GPU that loads the library until exit): the call makes no cudaMalloc / cudaFree / allocating library call.
 * nsp > ECO_NOALLOC_MAX_SPECIES is rejected (non-zero return, eco_last_error says so).  `out` (nsp int64) may be host or
 * device memory.  All calls on one GPU share the array -- the used prefix is cleared before each launch -- so calls that
 * overlap on the GPU (two models on one device driven concurrently) race; serialise them. */
#define ECO_NOALLOC_MAX_SPECIES 16384
int eco_species_totals_noalloc(eco_model* m, int64_t* out, size_t nsp);
/* Model<M>::checkpoint / restart, include/ecolab.h:88-94 */
int eco_checkpoint(eco_model* m, const char* file);
int eco_restart(eco_model* m, const char* file);

/* ---- building blocks for the multi-GPU driver (one process per GPU) ----------------
 * The collectives themselves (halo rows, alive-mask all-reduce, mutant all-gather)
 * run in the host language over NCCL; these calls are the per-rank halves. */
/* migrate on the owned slab using the ghost rows (or peer rows) as neighbours;
 * *abs_delta_sum = sum |delta| over owned cells, NOT halved. (:631-664) */
int eco_migrate_local(eco_model* m, uint64_t* abs_delta_sum);
/* the same without waiting for the kernel: *total_dev = device address of the uint64 sum |delta| (valid on the model's
 * stream after the launch), so the all-reduce of SpatialModel::migrate's return value (:667-672) can be enqueued
 * behind the kernel without a host round trip */
int eco_migrate_local_async(eco_model* m, void** total_dev);
/* alive mask of eco_alive_mask left ON THE DEVICE: *mask_dev = device address of nsp bytes (valid on the model's stream
 * after the launch), so the logical-or all-reduce of SpatialModel::condense (MPI_LOR, :182-186) runs on it in place, without a
 * host round trip before the collective; the reduced mask is then read once and handed to eco_condense_apply */
int eco_alive_mask_device(eco_model* m, void** mask_dev, size_t nsp);
/* device pointers + byte size of the two boundary rows to send and the two ghost rows to fill */
int eco_halo_ptrs(eco_model* m, void** send_top, void** send_bottom, void** recv_top, void** recv_bottom,
                  size_t* row_bytes);
/* Halo over peer memory instead of a ghost-row exchange (replaces graphcode's prepareNeighbours,
 * graphcode/prepare_neighbours.cc:17-72, on NVLink): every rank exports an opaque blob (CUDA-IPC handles of its two
 * density buffers and of its epoch counter + its slab geometry), the host language hands each rank the blobs of the ranks
 * owning the rows above (lower y) and below its slab (NULL = none), and from then on eco_migrate_local's boundary units
 * read the neighbours' boundary rows directly, ordered by device-side epoch counters (no collective, no host round trip).
 * All ranks must then call migrate the same number of times.  A (re)allocation of the density buffers (scalar
 * "buf_version" changes) disconnects; export and connect again. */
#define ECO_PEER_BLOB_BYTES 256
int eco_peer_export(eco_model* m, void* blob, size_t blob_bytes);
int eco_peer_connect(eco_model* m, const void* top_blob, const void* bottom_blob);
int eco_peer_disconnect(eco_model* m);
/* condense halves: local alive mask (:173-179); apply a global mask (:188-194) */
int eco_alive_mask(eco_model* m, uint8_t* alive, size_t nsp);
int eco_condense_apply(eco_model* m, const uint8_t* alive, size_t nsp, unsigned* removed);
/* mutate halves: per-cell speciation on the owned cells (:224-227, EcolabPoint::mutate :278-314),
 * records sorted by (cell, slot); then ModelData::mutate (:362-470) on the gathered parent list;
 * then the density append (:272-275). */
int eco_mutate_cells(eco_model* m, size_t* nrecords);                      /* runs the kernel, keeps the records */
int eco_mutate_records(eco_model* m, eco_mut_record* out, size_t cap);      /* copies them out (cap >= nrecords) */
int eco_mutate_model(eco_model* m, const uint32_t* parents, size_t n);
int eco_append_species(eco_model* m, const uint32_t* birth_cells, size_t n);

/* ---- module-level state of the reference's `ecolab` module --------------------------
 * array_urand.seed (include/random_basic.h:15: srand) and the process-wide
 * array_grand running-CLT generator (src/random.cc:59-64, src/arrays.cc:24-25). */
void eco_array_urand_seed(int seed);
/* array_urand.rand(): n values of rand()/RAND_MAX from the same stream (src/random.cc:57, include/random_basic.h:14) */
void eco_array_urand_rand(double* out, size_t n);
void eco_array_grand_reset(void);

/* tuning / profiling knobs (0 = automatic): "gen_dual" (-1: one species-row block per consumer warp of generate(1)),
 * "gen_stages", "gen_nobsmem" (1: beta through L1/L2 even where it fits in shared memory), "mig_rows" (grid rows per work
 * unit of migrate), "gen_dense" (1 = tensor-core
 * generate for dense beta; never automatic), "dense_slices", "timing" (1 = record a CUDA-event pair around every
 * generate/migrate launch, 0 = off and reset) */
int eco_set_tuning(eco_model* m, const char* knob, int value);
/* device time per kernel kind since timing was switched on: out = {generate_ms, generate_launches,
 * migrate_ms, migrate_launches, ...}; *n = number of doubles available */
int eco_kernel_times(eco_model* m, double* out, size_t cap, size_t* n);
/* x = n + n*(r + beta.n) of ONE generate sub-step BEFORE EcolabPoint::ROUND (:99-102), cell-major
 * [cell][nsp] for the owned cells; densities and tstep are not changed (float-tolerance check) */
int eco_pre_rounding(eco_model* m, double* out, size_t ncells, size_t nsp);

#ifdef __cplusplus
}
#endif
#endif /* ECOLAB_B200_H */
