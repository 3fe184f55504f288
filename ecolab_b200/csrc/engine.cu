// engine.cu -- device store + kernel launches for the EcoLab population update on sm_100a.
// Kernels: k_generate.cuh (Lotka-Volterra step), k_migrate.cuh (migration), k_misc.cuh (condense,
// mutate, observables, layout conversion); layout / RNG contract: k_common.cuh.
// Reference semantics (cited per kernel) come from models/ecolab_model.cc of
// highperformancecoder/ecolab; this file shares no code with it.
#include "engine.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>

#include "k_common.cuh"
#include "k_generate.cuh"
#include "k_migrate.cuh"
#include "k_misc.cuh"
#include "k_dense.cuh"

#define CK(x) ::eco::check_cuda((int)(x), #x, __FILE__, __LINE__)
#define ST ((cudaStream_t)stream_)

namespace eco {

void check_cuda(int err, const char* what, const char* file, int line) {
  if (err != (int)cudaSuccess) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", err,
             cudaGetErrorString((cudaError_t)err), file, line, what);
    throw std::runtime_error(buf);
  }
}

namespace {

unsigned long long* noalloc_totals_ptr(int device);   // engine_ops.inc: address of the static totals scratch on `device`

constexpr size_t kSmemPerSm = 227 * 1024;
inline uint32_t round_up32(uint32_t n) { return (n + 31u) & ~31u; }

// Tile width for a species capacity: the widest tile of which two fit the generate pipeline's 128 KB ring
// (the rest of the SM's 256 KB stays L1 and holds beta); failing that the widest that fits the SM twice, then once.
uint32_t pick_tile_width(uint32_t cap) {
  for (uint32_t tw : {32u, 16u, 8u, 4u})
    if (2 * (size_t)cap * tw * 4 <= 128 * 1024) {
      // above 1024 species beta never fits beside the tiles and is read through L2 whatever L1 is left: take the next wider
      // tile if two of them still fit the SM (1025-1784 species: 16 cells, 2049-3568: 8).  The narrow tiles are the slow ones
      // (4 cells: 32 species rows per warp access, every gather a 16-byte piece of a different row) -- measured on the
      // scripts' cadence at 64 x 64 cells, 1700-3300 species: generate(100) 20-26 ms with 4-cell tiles.
      if (tw <= 8 && 2 * (size_t)cap * (2 * tw) * 4 + 4096 <= kSmemPerSm) return 2 * tw;
      return tw;
    }
  if (2 * ((size_t)cap * 16 + 64) + (size_t)cap + 1024 <= kSmemPerSm) return 4;
  if ((size_t)cap * 16 + (size_t)cap + 1024 <= kSmemPerSm) return 4;
  throw std::runtime_error("species count exceeds the shared-memory tile (nsp*16 B must fit in 227 KB)");
}

// Capacity (species rows allocated) for nsp species: 25 % head-room so that speciation does not re-layout every round, but
// never so much that the head-room alone costs a tile width -- the kernels stage nsp rows, the tile width is chosen for the
// capacity (scripts' cadence at 64 x 64 cells: 3289 species under a 4288-row capacity ran generate(100) on 4-cell tiles).
uint32_t pick_capacity(uint32_t nsp) {
  const uint32_t cap = round_up32(nsp + nsp / 4 + 16);
  for (uint32_t lim : {512u, 1760u, 3552u})   // largest capacities (multiples of 32) of the 32-, 16- and 8-cell tiles
    if (nsp <= lim && cap > lim) return lim;
  return cap;
}

}  // namespace

// ===========================================================================================
// lifetime / storage
// ===========================================================================================
Engine::Engine(int device) : device_(device) {
  CK(cudaSetDevice(device_));
  cudaStream_t s;
  CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  stream_ = s;
  own_stream_ = true;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device_));
  num_sms_ = prop.multiProcessorCount;
  max_smem_ = (int)prop.sharedMemPerBlockOptin;
  CK(cudaMalloc(&d_total_, 4 * sizeof(unsigned long long)));
  CK(cudaMemset(d_total_, 0, 4 * sizeof(unsigned long long)));
  CK(cudaMallocHost(&h_total_, 4 * sizeof(unsigned long long)));
  memset(h_total_, 0, 4 * sizeof(unsigned long long));
  CK(cudaMalloc(&d_flags_, 4 * sizeof(unsigned int)));
  CK(cudaMemset(d_flags_, 0, 4 * sizeof(unsigned int)));
  cudaEvent_t ev;
  CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  ev_total_ = ev;
  noalloc_totals_ptr(device_);   // set-up half of species_totals_noalloc(): resolve the static scratch once per device
}

Engine::~Engine() {
  cudaSetDevice(device_);
  cudaDeviceSynchronize();
  try { peer_disconnect(); } catch (...) {}
  free_buffers();
  cudaFree(d_flags_);
  cudaFree(d_boff_); cudaFree(d_stream_); cudaFree(d_csr_);
  if (h_csr_) cudaFreeHost(h_csr_);
  cudaFree(d_aop_); cudaFree(d_bop_); cudaFree(d_dsp_); cudaFree(d_partial_); cudaFree(d_maxn_); cudaFree(dense_prof_);
  if (h_maxn_) cudaFreeHost(h_maxn_);
  cudaFree(d_vec_); cudaFree(d_mask_); cudaFree(d_map_); cudaFree(d_total_); cudaFree(d_rec_); cudaFree(d_scan_); cudaFree(d_stage_); cudaFree(d_scratch_);
  for (auto& p : d_band_) cudaFree(p);
  for (auto& e : ev_band_) if (e) cudaEventDestroy((cudaEvent_t)e);
  if (st_in_) cudaStreamDestroy((cudaStream_t)st_in_);
  if (st_out_) cudaStreamDestroy((cudaStream_t)st_out_);
  cudaFreeHost(h_total_);
  if (ev_total_) cudaEventDestroy((cudaEvent_t)ev_total_);
  for (auto& e : ev_pending_) { cudaEventDestroy((cudaEvent_t)e.start); cudaEventDestroy((cudaEvent_t)e.stop); }
  for (void* e : ev_pool_) cudaEventDestroy((cudaEvent_t)e);
  if (own_stream_ && stream_) cudaStreamDestroy(ST);
}

void Engine::set_stream(void* s) {
  CK(cudaSetDevice(device_));
  if (stream_) CK(cudaStreamSynchronize(ST));
  if (own_stream_ && stream_) cudaStreamDestroy(ST);
  stream_ = s;
  own_stream_ = false;
}

void Engine::sync() {
  CK(cudaSetDevice(device_));
  CK(cudaStreamSynchronize(ST));
}

void Engine::free_buffers() {
  peer_disconnect();   // the neighbours' row pointers were computed for the old geometry
  ++buf_version_;
  for (int i = 0; i < 2; ++i) {
    if (buf_[i]) cudaFree(buf_[i]);
    buf_[i] = nullptr;
  }
}

size_t Engine::bytes_allocated() const {
  const size_t per = (size_t)(ny_ + 2) * row_stride() * sizeof(int32_t);
  return (buf_[0] ? per : 0) + (buf_[1] ? per : 0);
}

int32_t* Engine::row0(int which) const { return buf_[which] + row_stride(); }

static Layout make_layout(int32_t* base, uint32_t tw, uint32_t cap, uint32_t ntx, uint32_t nx, uint32_t ny, uint32_t y0,
                          uint32_t nyg, uint32_t nsp) {
  Layout L;
  L.base = base;
  L.tile_stride = (size_t)cap * tw;
  L.row_stride = (size_t)ntx * cap * tw;
  L.tw = tw; L.ntx = ntx; L.nx = nx; L.ny = ny; L.y0 = y0; L.nyg = nyg; L.nsp = nsp; L.cap = cap;
  return L;
}
#define LAYOUT(which) make_layout(row0(which), tw_, cap_, ntx_, nx_, ny_, y0_, nyg_, nsp_)

void Engine::alloc_buffers(uint32_t cap) {
  free_buffers();
  cap_ = cap;
  cur_ = 0;
  if (cap_ == 0) { tw_ = 32; ntx_ = (nx_ + tw_ - 1) / tw_; return; }
  tw_ = pick_tile_width(cap_);
  ntx_ = (nx_ + tw_ - 1) / tw_;
  const size_t bytes = (size_t)(ny_ + 2) * row_stride() * sizeof(int32_t);
  CK(cudaMalloc(&buf_[0], bytes));
  CK(cudaMemsetAsync(buf_[0], 0, bytes, ST));
}

void Engine::ensure_other() {
  if (buf_[cur_ ^ 1]) return;
  const size_t bytes = (size_t)(ny_ + 2) * row_stride() * sizeof(int32_t);
  CK(cudaMalloc(&buf_[cur_ ^ 1], bytes));
  CK(cudaMemsetAsync(buf_[cur_ ^ 1], 0, bytes, ST));
  ++buf_version_;
}

void Engine::set_grid(uint32_t nx, uint32_t ny_global, uint32_t y0, uint32_t ny, uint32_t nsp) {
  CK(cudaSetDevice(device_));
  sync();
  nx_ = nx; nyg_ = ny_global; y0_ = y0; ny_ = ny;
  nsp_ = nsp;
  alloc_buffers(round_up32(nsp));
}

void Engine::resize_species(uint32_t nsp) {
  CK(cudaSetDevice(device_));
  if (nsp <= cap_) {
    if (nsp > nsp_ && buf_[cur_]) {   // re-zero the rows being brought to life (ghost rows too)
      bounds_sync();
      const uint32_t ntiles_all = ntx_ * (ny_ + 2);
      zero_rows_kernel<<<(unsigned)std::min<unsigned long long>(4096, (unsigned long long)ntiles_all * (nsp - nsp_) * tw_ / 256 + 1),
                         256, 0, ST>>>(LAYOUT(cur_), ntiles_all, buf_[cur_], nsp_, nsp);
      CK(cudaGetLastError());
      ++launches_;
    }
    nsp_ = nsp;
    return;
  }
  // grow: new capacity with head-room (and possibly a narrower tile), re-layout into a fresh buffer
  relayout(pick_capacity(nsp), nsp);
}

// After species were removed: if the survivors fit a wider tile, re-layout to it (one pass over the densities, paid back by
// the first generate()).  The decision depends on the species count and the GLOBAL grid only, so every rank of a
// partitioned grid takes it alike (the halo rows are exchanged in layout order).
void Engine::retile_after_shrink() {
  if (!buf_[cur_] || !nsp_) return;
  const uint32_t want_cap = pick_capacity(nsp_);
  if (want_cap >= cap_ || pick_tile_width(want_cap) <= tw_) return;
  if ((size_t)nx_ * nyg_ * want_cap * sizeof(int32_t) > ((size_t)16 << 30)) return;   // needs a second buffer: small grids only
  CK(cudaSetDevice(device_));
  relayout(want_cap, nsp_);
}

// move the densities into a fresh buffer of capacity new_cap (tile width chosen for it) holding nsp species
void Engine::relayout(uint32_t new_cap, uint32_t nsp) {
  const uint32_t new_tw = pick_tile_width(new_cap);
  peer_disconnect();
  ++buf_version_;
  int32_t* old_cur = buf_[cur_];
  int32_t* old_other = buf_[cur_ ^ 1];
  CK(cudaStreamSynchronize(ST));
  if (old_other) CK(cudaFree(old_other));
  buf_[0] = buf_[1] = nullptr;
  const Layout S = make_layout(old_cur, tw_, cap_, ntx_, nx_, ny_ + 2, y0_, nyg_, nsp_);
  const uint32_t old_nsp = nsp_;
  cap_ = new_cap; tw_ = new_tw; ntx_ = (nx_ + tw_ - 1) / tw_;
  const size_t bytes = (size_t)(ny_ + 2) * row_stride() * sizeof(int32_t);
  int32_t* nb = nullptr;
  CK(cudaMalloc(&nb, bytes));
  CK(cudaMemsetAsync(nb, 0, bytes, ST));
  buf_[0] = nb;
  cur_ = 0;
  bounds_sync(old_cur, (size_t)(ny_ + 2) * S.row_stride * sizeof(int32_t));   // the new buffer + the one being read
  if (old_cur && old_nsp) {
    const Layout D = make_layout(nb, tw_, cap_, ntx_, nx_, ny_ + 2, y0_, nyg_, nsp);
    const unsigned long long n = (unsigned long long)(ny_ + 2) * old_nsp * std::min(S.ntx * S.tw, D.ntx * D.tw);
    relayout_kernel<<<(unsigned)std::min<unsigned long long>(148 * 16, n / 256 + 1), 256, 0, ST>>>(S, D, ny_ + 2, old_nsp);
    CK(cudaGetLastError());
    ++launches_;
  }
  CK(cudaStreamSynchronize(ST));
  if (old_cur) CK(cudaFree(old_cur));
  buf_[0] = nb;
  cur_ = 0;
  nsp_ = nsp;
}

// ===========================================================================================
// density access
// ===========================================================================================
void Engine::set_density(size_t cell, const int32_t* d, size_t n) {
  CK(cudaSetDevice(device_));
  if (n > nsp_) throw std::runtime_error("set_density: more species than the model has");
  if (cell >= ncells_local()) throw std::runtime_error("set_density: cell out of range");
  const uint32_t yl = (uint32_t)(cell / nx_), x = (uint32_t)(cell % nx_);
  int32_t* dst = row0(cur_) + ((size_t)yl * ntx_ + x / tw_) * tile_stride() + (x % tw_);
  if (n) CK(cudaMemcpy2DAsync(dst, tw_ * 4, d, 4, 4, n, cudaMemcpyHostToDevice, ST));
  if (n < nsp_) CK(cudaMemset2DAsync(dst + n * tw_, tw_ * 4, 0, 4, nsp_ - n, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::get_density(size_t cell, int32_t* d, size_t n) {
  CK(cudaSetDevice(device_));
  if (n > nsp_) throw std::runtime_error("get_density: more species than the model has");
  if (cell >= ncells_local()) throw std::runtime_error("get_density: cell out of range");
  if (!n) return;
  const uint32_t yl = (uint32_t)(cell / nx_), x = (uint32_t)(cell % nx_);
  const int32_t* src = row0(cur_) + ((size_t)yl * ntx_ + x / tw_) * tile_stride() + (x % tw_);
  CK(cudaMemcpy2DAsync(d, 4, src, tw_ * 4, 4, n, cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::set_density_all_dev(const int32_t* dev, size_t nsp) {
  CK(cudaSetDevice(device_));
  bounds_sync();
  if (nsp != nsp_) throw std::runtime_error("set_density_all: species count mismatch");
  const unsigned long long nwork = (unsigned long long)((nx_ + 31) / 32) * ny_ * ((nsp_ + 31) / 32);
  if (!nwork) return;
  transpose_kernel<true><<<(unsigned)std::min<unsigned long long>(nwork, 148ull * 64), 1024, 0, ST>>>(
      LAYOUT(cur_), const_cast<int32_t*>(dev));
  CK(cudaGetLastError());
  ++launches_;
}

void Engine::get_density_all_dev(int32_t* dev, size_t nsp) {
  CK(cudaSetDevice(device_));
  bounds_sync();
  if (nsp != nsp_) throw std::runtime_error("get_density_all: species count mismatch");
  const unsigned long long nwork = (unsigned long long)((nx_ + 31) / 32) * ny_ * ((nsp_ + 31) / 32);
  if (!nwork) return;
  transpose_kernel<false><<<(unsigned)std::min<unsigned long long>(nwork, 148ull * 64), 1024, 0, ST>>>(LAYOUT(cur_), dev);
  CK(cudaGetLastError());
  ++launches_;
}

void Engine::set_density_all(const int32_t* host, size_t nsp) {
  CK(cudaSetDevice(device_));
  const size_t n = ncells_local() * nsp;
  if (n > stage_cap_) {
    if (d_stage_) CK(cudaFree(d_stage_));
    CK(cudaMalloc(&d_stage_, n * sizeof(int32_t)));
    stage_cap_ = n;
  }
  CK(cudaMemcpyAsync(d_stage_, host, n * sizeof(int32_t), cudaMemcpyHostToDevice, ST));
  set_density_all_dev(d_stage_, nsp);
  CK(cudaStreamSynchronize(ST));
}

void Engine::get_density_all(int32_t* host, size_t nsp) {
  CK(cudaSetDevice(device_));
  const size_t n = ncells_local() * nsp;
  if (n > stage_cap_) {
    if (d_stage_) CK(cudaFree(d_stage_));
    CK(cudaMalloc(&d_stage_, n * sizeof(int32_t)));
    stage_cap_ = n;
  }
  get_density_all_dev(d_stage_, nsp);
  CK(cudaMemcpyAsync(host, d_stage_, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::get_density_rows(uint32_t r0, uint32_t nrows, int32_t* host, size_t nsp) {
  CK(cudaSetDevice(device_));
  bounds_sync();
  if (nsp != nsp_) throw std::runtime_error("get_density_rows: species count mismatch");
  if ((uint64_t)r0 + nrows > ny_) throw std::runtime_error("get_density_rows: rows outside the owned slab");
  const size_t n = (size_t)nx_ * nrows * nsp;
  if (!n) return;
  if (n > stage_cap_) {
    if (d_stage_) CK(cudaFree(d_stage_));
    CK(cudaMalloc(&d_stage_, n * sizeof(int32_t)));
    stage_cap_ = n;
  }
  const unsigned long long nwork = (unsigned long long)((nx_ + 31) / 32) * nrows * ((nsp_ + 31) / 32);
  transpose_kernel<false><<<(unsigned)std::min<unsigned long long>(nwork, 148ull * 64), 1024, 0, ST>>>(
      make_layout(row0(cur_) + (size_t)r0 * row_stride(), tw_, cap_, ntx_, nx_, nrows, y0_ + r0, nyg_, nsp_), d_stage_);
  CK(cudaGetLastError());
  ++launches_;
  CK(cudaMemcpyAsync(host, d_stage_, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ST));
  CK(cudaStreamSynchronize(ST));
}

void Engine::set_density_rows(uint32_t r0, uint32_t nrows, const int32_t* host, size_t nsp) {
  CK(cudaSetDevice(device_));
  bounds_sync();
  if (nsp != nsp_) throw std::runtime_error("set_density_rows: species count mismatch");
  if ((uint64_t)r0 + nrows > ny_) throw std::runtime_error("set_density_rows: rows outside the owned slab");
  const size_t n = (size_t)nx_ * nrows * nsp;
  if (!n) return;
  if (n > stage_cap_) {
    if (d_stage_) CK(cudaFree(d_stage_));
    CK(cudaMalloc(&d_stage_, n * sizeof(int32_t)));
    stage_cap_ = n;
  }
  CK(cudaMemcpyAsync(d_stage_, host, n * sizeof(int32_t), cudaMemcpyHostToDevice, ST));
  const unsigned long long nwork = (unsigned long long)((nx_ + 31) / 32) * nrows * ((nsp_ + 31) / 32);
  transpose_kernel<true><<<(unsigned)std::min<unsigned long long>(nwork, 148ull * 64), 1024, 0, ST>>>(
      make_layout(row0(cur_) + (size_t)r0 * row_stride(), tw_, cap_, ntx_, nx_, nrows, y0_ + r0, nyg_, nsp_), d_stage_);
  CK(cudaGetLastError());
  ++launches_;
  CK(cudaStreamSynchronize(ST));
}

void Engine::fill_density(int32_t v) {
  CK(cudaSetDevice(device_));
  bounds_sync();
  const uint32_t ntiles = ntx_ * ny_;
  if (!ntiles || !nsp_) return;
  fill_kernel<<<(unsigned)std::min<unsigned long long>(148 * 32, (unsigned long long)ntiles * nsp_ * tw_ / 256 + 1), 256, 0, ST>>>(
      LAYOUT(cur_), v);
  CK(cudaGetLastError());
  ++launches_;
}

void Engine::fill_density_random(uint32_t seed, int32_t lo, int32_t hi) {
  CK(cudaSetDevice(device_));
  bounds_sync();
  const uint32_t ntiles = ntx_ * ny_;
  if (!ntiles || !nsp_) return;
  const unsigned long long n = (unsigned long long)ntiles * nsp_ * (tw_ / 4);
  fill_random_kernel<<<(unsigned)std::min<unsigned long long>(148 * 32, n / 256 + 1), 256, 0, ST>>>(
      LAYOUT(cur_), seed, lo, (uint32_t)(hi - lo) + 1u);
  CK(cudaGetLastError());
  ++launches_;
}

#include "engine_ops.inc"
#include "engine_dense.inc"

}  // namespace eco
