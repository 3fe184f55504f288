"""spatial_ecolab: the grid model (SpatialModel, models/ecolab_model.h:128-151) and its
multi-GPU driver.

One process per GPU.  The nx x ny cell grid (cell id = x + nx*y, makeId,
models/ecolab_model.h:137) is split into contiguous row slabs, rank g owning rows
[g*ny/G, (g+1)*ny/G); this replaces graphcode's round-robin / ParMETIS partition
(models/ecolab_model.cc:591, graphcode/partition.cc) and its MPI object exchange.
Species parameters and beta are replicated.  Communication happens only where the
reference communicates:

  migrate   halo rows (one grid row of densities per side) -- the reference's
            prepareNeighbours (graphcode/prepare_neighbours.cc:17-72).  Default on
            GPUs: the migrate kernel's boundary units read the neighbour ranks'
            boundary rows directly out of their buffers over NVLink (CUDA IPC
            mappings, device-side epoch counters, no collective; eco_peer_*);
            fallback (ECOLAB_B200_HALO=nccl, or IPC unavailable): NCCL send/recv
            into ghost rows.  Plus a sum all-reduce of the migration total (:667-672)
  condense  max all-reduce of the alive mask (MPI_LOR, :182-186)
  mutate    all-gather of the (cell, parent slot, count) records (:250), rank 0
            generates the new species' parameters and broadcasts only the appended
            tails (the reference broadcasts the whole ModelData, :262-263)

The collectives go through ``torch.distributed`` (NCCL on GPUs; gloo works for the
host-side logic and is what the CPU tests use).
"""
import ctypes as C

import numpy as np

from . import _cabi
from ._cabi import EcolabError, MutRecord, check
from .model import _SCALARS_B, _SCALARS_F, EcoList, EcolabModelBase, _Cell, _Member, _ptr


def slab_bounds(ny, world, rank):
    """Rows [y0, y1) owned by `rank` of `world` (balanced, contiguous, ascending)."""
    return (rank * ny) // world, ((rank + 1) * ny) // world


def merge_records(per_rank):
    """Concatenate per-rank record arrays (each sorted by (cell, slot)) into the global
    order: ascending cell id, then slot.  Slabs are contiguous ascending cell ranges, so
    rank order is already global order; a stable sort keeps this true for any partition."""
    parts = [np.asarray(p, dtype=np.uint32).reshape(-1, 3) for p in per_rank]
    parts = [p for p in parts if len(p)]
    if not parts:
        return np.zeros((0, 3), dtype=np.uint32)
    rec = np.concatenate(parts)
    order = np.lexsort((rec[:, 1], rec[:, 0]))
    return rec[order]


def expand_records(rec):
    """gen_index (include/arrays.h:2477-2486): parent slot / birth cell repeated `count` times."""
    rec = np.asarray(rec, dtype=np.uint32).reshape(-1, 3)
    cnt = rec[:, 2].astype(np.int64)
    return np.repeat(rec[:, 1], cnt).astype(np.uint32), np.repeat(rec[:, 0], cnt).astype(np.uint32)


class Comm:
    """The four exchanges of the distributed model over torch.distributed."""

    def __init__(self, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.backend = dist.get_backend(group)
        self.dev = torch.device("cuda", torch.cuda.current_device()) if self.backend == "nccl" else torch.device("cpu")

    def _t(self, a):
        return self.torch.from_numpy(np.ascontiguousarray(a)).to(self.dev)

    def allreduce_max_u8(self, mask):
        t = self._t(mask.astype(np.uint8))
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return t.cpu().numpy()

    def allreduce_sum(self, v):
        t = self._t(np.array([v], dtype=np.int64))
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return int(t.cpu()[0])

    def allgather_records(self, rec):
        rec = np.asarray(rec, dtype=np.uint32).reshape(-1, 3)
        n = self._t(np.array([len(rec)], dtype=np.int64))
        counts = [self.torch.zeros_like(n) for _ in range(self.world)]
        self.dist.all_gather(counts, n, group=self.group)
        counts = [int(c.cpu()[0]) for c in counts]
        mx = max(counts)
        if mx == 0:
            return [np.zeros((0, 3), dtype=np.uint32)] * self.world
        pad = np.zeros((mx, 3), dtype=np.int32)
        pad[: len(rec)] = rec.view(np.int32)
        t = self._t(pad)
        outs = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(outs, t, group=self.group)
        return [o.cpu().numpy().view(np.uint32)[:c] for o, c in zip(outs, counts)]

    def bcast_arrays(self, arrays, src=0):
        """Broadcast a (possibly nested) dict of numpy arrays / scalars from `src` as an object (setup-time replica
        synchronisation: sizes are not known on the receivers)."""
        obj = [arrays if self.rank == src else None]
        self.dist.broadcast_object_list(obj, src=src, group=self.group)
        return obj[0]

    def bcast_tails(self, tails, f64_keys, u32_keys, src=0):
        """Broadcast the arrays appended by ModelData::mutate (a dict name -> 1-D array, plus "species" int32) from `src`
        as three tensors: one int64 header with the lengths, one float64 buffer, one int32 buffer -- no pickling, and on
        NCCL nothing but the header is read before the payload is on every device."""
        keys = list(f64_keys) + list(u32_keys) + ["species"]
        hdr = self._t(np.array([len(tails[k]) for k in keys] if self.rank == src else [0] * len(keys), dtype=np.int64))
        self.dist.broadcast(hdr, src=src, group=self.group)
        lens = [int(v) for v in hdr.cpu().numpy()]
        nf, ni = sum(lens[: len(f64_keys)]), sum(lens[len(f64_keys):])
        if self.rank == src:
            fb = np.concatenate([np.asarray(tails[k], dtype=np.float64) for k in f64_keys]) if nf else np.zeros(0)
            ib = np.concatenate([np.asarray(tails[k]).astype(np.uint32).view(np.int32) for k in u32_keys] +
                                [np.asarray(tails["species"], dtype=np.int32)]) if ni else np.zeros(0, dtype=np.int32)
        else:
            fb, ib = np.zeros(nf, dtype=np.float64), np.zeros(ni, dtype=np.int32)
        tf, ti = self._t(fb), self._t(ib)
        if nf:
            self.dist.broadcast(tf, src=src, group=self.group)
        if ni:
            self.dist.broadcast(ti, src=src, group=self.group)
        if self.rank == src:
            return tails
        fb, ib = tf.cpu().numpy(), ti.cpu().numpy()
        out, of, oi = {}, 0, 0
        for k, n in zip(keys, lens):
            if k in f64_keys:
                out[k] = fb[of:of + n].copy(); of += n
            elif k == "species":
                out[k] = ib[oi:oi + n].copy(); oi += n
            else:
                out[k] = ib[oi:oi + n].view(np.uint32).copy(); oi += n
        return out

    def sendrecv_rows(self, send_up, recv_up, send_down, recv_down):
        """Exchange boundary rows with rank-1 ("up", lower y) and rank+1 ("down").  Tensors alias device memory."""
        ops = []
        P2P = self.dist.P2POp
        if self.rank > 0:
            ops += [P2P(self.dist.isend, send_up, self.rank - 1, self.group), P2P(self.dist.irecv, recv_up, self.rank - 1, self.group)]
        if self.rank + 1 < self.world:
            ops += [P2P(self.dist.isend, send_down, self.rank + 1, self.group), P2P(self.dist.irecv, recv_down, self.rank + 1, self.group)]
        if ops:
            for r in self.dist.batch_isend_irecv(ops):
                r.wait()


# ---------------------------------------------------------------------------------
# The distributed algorithms, written against the per-rank halves ("ops") so that the
# same code drives the CUDA engine (SpatialModel below) and, in the CPU tests, a
# numpy/oracle emulation of one rank under gloo.
#   ops.alive_mask() -> u8[nsp]            ops.condense_apply(mask) -> removed
#   ops.mutate_cells() -> u32[n,3]         ops.mutate_model(parents)
#   ops.lengths() / ops.tails(before) / ops.append_tails(tails)
#   ops.append_species(cells)              ops.migrate_local() -> sum|delta| (not halved)
#   ops.halo_tensors() -> (send_up, recv_up, send_down, recv_down) or None; ops.halo_done()
# ---------------------------------------------------------------------------------
def dist_migrate(ops, comm, wait=True):
    """SpatialModel::migrate (models/ecolab_model.cc:616-703) on a row-slab partition.  wait=False (device path only):
    nothing is read back; the caller fetches the total later with ops.last_migration()."""
    t = ops.halo_tensors()
    if t is not None:
        comm.sendrecv_rows(*t)
        ops.halo_done()
    dev_total = ops.migrate_local_device() if comm.backend == "nccl" and hasattr(ops, "migrate_local_device") else None
    if dev_total is not None:   # all-reduce enqueued behind the kernel: at most one host synchronisation per migrate, at the end
        if not wait:
            return None
        comm.dist.all_reduce(dev_total, op=comm.dist.ReduceOp.SUM, group=comm.group)
        return int(dev_total.cpu()[0]) // 2
    return comm.allreduce_sum(ops.migrate_local()) // 2   # :667-672, :702


def dist_condense(ops, comm):
    """SpatialModel::condense (:171-196): local mask, logical-or all-reduce, identical compaction everywhere.  On GPUs the
    mask never leaves the device before the collective: the all-reduce runs in place on the kernel's output and the reduced
    mask is read once (ModelData::condense needs it on the host)."""
    dev = ops.alive_mask_device() if comm.backend == "nccl" and hasattr(ops, "alive_mask_device") else None
    if dev is not None:
        comm.dist.all_reduce(dev, op=comm.dist.ReduceOp.MAX, group=comm.group)
        return ops.condense_apply(np.ascontiguousarray(dev.cpu().numpy()))
    mask = np.ascontiguousarray(comm.allreduce_max_u8(ops.alive_mask()))
    return ops.condense_apply(mask)


def dist_mutate(ops, comm):
    """SpatialModel::mutate (:215-276): per-cell speciation, gather of the records, parameter
    generation on rank 0, broadcast of the appended tails only, density append everywhere."""
    merged = merge_records(comm.allgather_records(ops.mutate_cells()))
    parents, cells = expand_records(merged)
    if len(parents) == 0:
        return 0
    before = ops.lengths()
    tails = None
    if comm.rank == 0:   # ModelData::mutate on rank 0 only (:251-258): it draws from rank 0's libc stream
        ops.mutate_model(parents)
        tails = ops.tails(before)
    tails = comm.bcast_tails(tails, ops._TAILS_F64, ops._TAILS_U32, src=0)
    if comm.rank != 0:
        ops.append_tails(tails)
    ops.append_species(cells)
    return len(parents)


class _DevMem:
    """Expose a raw device range to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, nbytes, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": (nbytes // int(typestr[2:]),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


class _OwnedCell(_Cell):
    """A cell of this rank's slab in a distributed run: reads are collective (see SpatialModel.cell)."""

    def __init__(self, model, cid, j):
        super().__init__(model, cid)
        self.density = _Member(lambda: model._bcast_density(cid, j), lambda v: model._set_density(cid, v), name=f"cell[{cid}].density")

    def nsp(self):
        return sum(1 for v in self.density() if v != 0)


class _RemoteCell:
    """A cell owned by another rank: assignments are dropped (its owner executes the same statement), reads are
    collective broadcasts from the owner."""

    def __init__(self, model, cid, j):
        self._m, self._id = model, cid

        def _set(v):
            a = np.asarray(list(v) if not isinstance(v, np.ndarray) else v)
            if (a < 0).any():
                raise EcolabError("density must be non-negative (reference asserts density>=0, models/ecolab_model.cc:108)")
            model._note_remote_species(len(a))

        self.density = _Member(lambda: model._bcast_density(cid, j), _set, name=f"cell[{cid}].density")
        self.salt = _Member(lambda: model._salt(cid), name="salt")
        self.id = _Member(lambda: cid, name="id")

    def nsp(self):
        return sum(1 for v in self.density() if v != 0)

    def _list(self):
        return ["density", "salt", "id"]


class SpatialModel(EcolabModelBase):
    """spatial_ecolab.  Single process: the whole grid on one GPU.  Under torch.distributed
    (world > 1): this rank's row slab, with the exchanges described in the module docstring."""

    _TAILS_F64 = ("create", "repro_rate", "mutation", "migration", "interaction.diag", "interaction.val")
    _TAILS_U32 = ("interaction.row", "interaction.col")

    def __init__(self, device=0, comm=None):
        super().__init__(device)
        self._comm = comm
        self._nx = self._ny = 1
        self._y0, self._y1 = 0, 1
        self._peer_version = None      # buf_version the neighbours' buffers were mapped for
        self._peer_failed = False
        self._replica_dirty = False    # a ModelData member was assigned since the replicas were last made identical

    # ---- replicated ModelData -----------------------------------------------------
    # Every rank holds the whole ModelData and runs the same script, but not every assignment is rank-independent: the
    # reference's scripts seed `array_urand` with 10+myid() ("different seed across processes", models/spatial_ecolab.py:14),
    # so random_interaction() draws a different beta on every rank, and an unseeded Python random() differs too.  The
    # reference makes the replicas identical again when SpatialModel::mutate broadcasts rank 0's ModelData (:262-263).  Here
    # rank 0's parameters are broadcast before the first hot-path call that follows an assignment, so a partitioned run is
    # the single-process run of the same script from the first generate on (tests/test_gpu_ref_scripts.py checks that).
    def _set_array(self, name, kind, v):
        super()._set_array(name, kind, v)
        self._replica_dirty = True

    def _set_scalar(self, name, v):
        super()._set_scalar(name, v)
        self._replica_dirty = True

    def random_interaction(self, conn, sigma):
        super().random_interaction(conn, sigma)
        self._replica_dirty = True

    def _sync_replicas(self):
        if self.world == 1 or not self._replica_dirty:
            return
        self._replica_dirty = False
        names = [(n, "f64") for n in self._TAILS_F64] + [(n, "u32") for n in self._TAILS_U32] + [("species", "i32")]
        scal = _SCALARS_F + _SCALARS_B
        blob = None
        if self.rank == 0:
            blob = {"arrays": {n: self._get_array(n, k) for n, k in names}, "scalars": {n: self._scalar(n) for n in scal}}
        blob = self._comm.bcast_arrays(blob, src=0)
        if self.rank != 0:
            for n, k in names:
                EcolabModelBase._set_array(self, n, k, blob["arrays"][n])
            for n in scal:
                EcolabModelBase._set_scalar(self, n, blob["scalars"][n])

    def generate(self, niter=1):
        self._sync_replicas()
        super().generate(niter)

    # ---- distribution ------------------------------------------------------------
    def attach(self, comm):
        """Join a process group; call before setGrid."""
        self._comm = comm

    @property
    def world(self):
        return self._comm.world if self._comm else 1

    @property
    def rank(self):
        return self._comm.rank if self._comm else 0

    def setGrid(self, nx, ny):
        nx, ny = int(nx), int(ny)
        self._nx, self._ny = nx, ny
        if self.world > 1:
            if ny < self.world:
                raise EcolabError(f"setGrid: {ny} grid rows cannot be split over {self.world} ranks")
            self._y0, self._y1 = slab_bounds(ny, self.world, self.rank)
            check(self._L.eco_set_scalar(self._h, b"nprocs", float(self.world)))
            check(self._L.eco_set_grid_slab(self._h, nx, ny, self._y0, self._y1 - self._y0))
        else:
            self._y0, self._y1 = 0, ny
            check(self._L.eco_set_grid(self._h, nx, ny))

    def restart(self, file):
        """Model<M>::restart (include/ecolab.h:92-94); the grid geometry comes back with the file."""
        super().restart(self._rank_file(file))
        self._nx, self._ny = int(self._scalar("numX")), int(self._scalar("numY"))
        self._y0 = int(self._scalar("y0"))
        self._y1 = self._y0 + int(self._scalar("ny_local"))
        self._peer_version = None

    def checkpoint(self, file):
        """One file per rank under torch.distributed (`file.rankN`): every rank holds its own slab."""
        super().checkpoint(self._rank_file(file))

    def _rank_file(self, file):
        return f"{file}.rank{self.rank}" if self.world > 1 else str(file)

    def owns(self, i, j):
        return self._y0 <= j < self._y1

    def cell(self, i, j):
        """EcolabCell of grid position (i, j).  Under torch.distributed every rank runs the same script, so the cell may
        belong to another rank: assignments to it are then accepted and dropped (the owner performs the same
        assignment), exactly as every MPI rank of the reference fills all cells before distributeObjects
        (graphcode/graphcode.h:475-482); reading it fetches the density from its owner (a collective: all ranks must
        make the same call, which an SPMD script does)."""
        cid = int(i) + self._nx * int(j)
        if self.world > 1 and not self.owns(i, j):
            return _RemoteCell(self, cid, int(j))
        if self.world > 1:
            return _OwnedCell(self, cid, int(j))
        return _Cell(self, cid)

    def cell_local(self, i, j):
        """The same for a cell this rank owns, without any collective (library code and tests; not the scripts' call)."""
        return _Cell(self, int(i) + self._nx * int(j))

    def _owner_of_row(self, j):
        for r in range(self.world):
            y0, y1 = slab_bounds(self._ny, self.world, r)
            if y0 <= j < y1:
                return r
        raise EcolabError("cell id outside the grid")

    def _bcast_density(self, cid, j):
        """Collective read of one cell's density from its owner."""
        import torch
        src = self._owner_of_row(j)
        ns = int(self._scalar("nsp"))
        a = np.asarray(self._get_density(cid), dtype=np.int32) if src == self.rank else np.zeros(ns, dtype=np.int32)
        t = torch.from_numpy(a).to(self._comm.dev)
        self._comm.dist.broadcast(t, src=src, group=self._comm.group)
        return EcoList(t.cpu().numpy().tolist())

    def nsp_local(self):
        """Living species per OWNED cell (ascending cell id)."""
        nc = int(self._scalar("ncells_local"))
        a = np.empty(nc, dtype=np.uint32)
        if nc:
            check(self._L.eco_nsp_per_cell(self._h, _ptr(a), nc))
        return a

    def nsp(self):
        """Living species per cell of the WHOLE grid, ascending cell id (SpatialModel::nsp, models/ecolab_model.cc:117-122:
        the reference gathers the per-rank counts with MPI); an all-gather under torch.distributed."""
        a = self.nsp_local()
        if self.world == 1:
            return EcoList(a.tolist())
        import torch
        sizes = [(slab_bounds(self._ny, self.world, r)[1] - slab_bounds(self._ny, self.world, r)[0]) * self._nx for r in range(self.world)]
        mx = max(sizes)
        pad = np.zeros(mx, dtype=np.int32)
        pad[: len(a)] = a.astype(np.int32)
        t = torch.from_numpy(pad).to(self._comm.dev)
        outs = [torch.zeros_like(t) for _ in range(self.world)]
        self._comm.dist.all_gather(outs, t, group=self._comm.group)
        return EcoList(np.concatenate([o.cpu().numpy()[:n] for o, n in zip(outs, sizes)]).tolist())

    # graphcode operations that have nothing left to do under a static slab partition: the partition is fixed by
    # setGrid, objects never move
    def partitionObjects(self):
        return None

    def distributeObjects(self):
        return None

    def gather(self):
        """Graph::gather (graphcode/gather.cc:17-38): afterwards rank 0 can read every cell.  Here every rank already
        reads any cell through its owner (cell(i,j).density() is a collective broadcast), so gathering is a
        synchronisation point: all kernels finished on every rank."""
        self.syncThreads()
        if self.world > 1:
            self._comm.dist.barrier(group=self._comm.group)

    def density_rows(self, y, nrows):
        """Whole grid rows [y, y+nrows) (all owned by this rank) as an int32 array [nrows*nx][nsp]."""
        ns = int(self._scalar("nsp"))
        a = np.empty((int(nrows) * self._nx, ns), dtype=np.int32)
        if a.size:
            check(self._L.eco_get_density_rows(self._h, int(y), int(nrows), _ptr(a), ns))
        return a

    def set_density_rows(self, y, a):
        a = np.ascontiguousarray(a, dtype=np.int32)
        check(self._L.eco_set_density_rows(self._h, int(y), a.shape[0] // self._nx, _ptr(a), a.shape[1]))

    # ---- halo over peer memory --------------------------------------------------------
    def _peers_wanted(self):
        import os
        return (self.world > 1 and self._comm.backend == "nccl" and not self._peer_failed
                and os.environ.get("ECOLAB_B200_HALO", "peer") != "nccl" and self._on_torch_stream())

    def _ensure_peers(self):
        """(Re)connect to the neighbour ranks' buffers when this rank's buffers were (re)allocated since the last
        exchange.  Buffer versions move in lock-step on all ranks (same species counts everywhere), so every rank takes
        this branch in the same migrate call; the exchange is a collective.  Returns True when the peer path is live."""
        if not self._peers_wanted() or not int(self._scalar("nsp")):
            return False
        if self._peer_version == int(self._scalar("buf_version")) and int(self._scalar("peer_connected")):
            return True
        import torch
        ok = 1
        blob = np.zeros(256, dtype=np.uint8)
        try:
            check(self._L.eco_peer_export(self._h, _ptr(blob), blob.size))
        except EcolabError:
            ok = 0
        t = torch.from_numpy(blob.copy()).to(self._comm.dev)
        outs = [torch.zeros_like(t) for _ in range(self.world)]
        self._comm.dist.all_gather(outs, t, group=self._comm.group)
        blobs = [o.cpu().numpy() for o in outs]
        if ok:
            top = blobs[self.rank - 1] if self.rank > 0 else None
            bot = blobs[self.rank + 1] if self.rank + 1 < self.world else None
            try:
                check(self._L.eco_peer_connect(self._h, _ptr(top) if top is not None else None, _ptr(bot) if bot is not None else None))
            except EcolabError:
                ok = 0
        flag = torch.tensor([ok], device=self._comm.dev)
        self._comm.dist.all_reduce(flag, op=self._comm.dist.ReduceOp.MIN, group=self._comm.group)   # all or nobody; also the barrier after connect
        if int(flag.item()) == 0:
            check(self._L.eco_peer_disconnect(self._h))
            self._peer_failed = True      # IPC not available here: NCCL ghost rows from now on
            return False
        self._peer_version = int(self._scalar("buf_version"))
        return True

    def _note_remote_species(self, n):
        """A density of n species was assigned to a cell of another rank: this rank's cells must have (at least) n
        species rows too, as SpatialModel::makeConsistent takes the maximum over all cells (:725-735)."""
        if n > int(self._scalar("nsp")) and self._y1 > self._y0:
            first = self._y0 * self._nx
            cur = np.asarray(self._get_density(first), dtype=np.int32)
            self._set_density(first, np.concatenate([cur, np.zeros(n - len(cur), dtype=np.int32)]))

    # ---- per-rank halves over the C ABI ("ops" protocol) -------------------------------
    def halo_tensors(self):
        import torch
        if not int(self._scalar("nsp")):
            return None
        if self._ensure_peers():
            return None          # the kernel reads the neighbours' rows itself
        st, sb, rt, rb = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        nb = C.c_size_t()
        check(self._L.eco_halo_ptrs(self._h, C.byref(st), C.byref(sb), C.byref(rt), C.byref(rb), C.byref(nb)))
        if nb.value == 0:
            return None
        dev = torch.device("cuda", self._device)
        T = lambda p: torch.as_tensor(_DevMem(p.value, nb.value), device=dev)   # noqa: E731
        if not self._on_torch_stream():
            self.syncThreads()   # generate ran on the engine's own stream; NCCL is ordered after torch's current stream
        return T(st), T(rt), T(sb), T(rb)

    def halo_done(self):
        import torch
        if not self._on_torch_stream():
            torch.cuda.current_stream(torch.device("cuda", self._device)).synchronize()

    def set_stream(self, torch_stream):
        """Run this model's kernels on a torch stream (eco_set_stream).  When it is also torch's CURRENT stream the
        multi-GPU driver needs no host synchronisation between kernels and collectives: NCCL orders itself after the
        current stream and the current stream waits for NCCL (work.wait())."""
        check(self._L.eco_set_stream(self._h, C.c_void_p(torch_stream.cuda_stream)))
        self._stream_handle = int(torch_stream.cuda_stream)

    def _on_torch_stream(self):
        import torch
        h = getattr(self, "_stream_handle", None)
        return h is not None and h == int(torch.cuda.current_stream(torch.device("cuda", self._device)).cuda_stream)

    def migrate_local_device(self):
        """migrate on the owned slab without waiting; returns the device-resident int64[1] sum |delta| (not halved), or
        None when the kernels do not run on torch's current stream (the caller then uses migrate_local)."""
        import torch
        if not self._on_torch_stream() or not int(self._scalar("nsp")):
            return None
        p = C.c_void_p()
        check(self._L.eco_migrate_local_async(self._h, C.byref(p)))
        return torch.as_tensor(_DevMem(p.value, 8, "<i8"), device=torch.device("cuda", self._device))

    def migrate_local(self):
        t = C.c_uint64()
        check(self._L.eco_migrate_local(self._h, C.byref(t)))
        return t.value

    def last_migration(self):
        """Total of the most recent migrate()/step(..., wait=False): blocks until it has finished (and, on a partitioned
        grid, sums over the ranks)."""
        t = C.c_uint64()
        check(self._L.eco_migrate_wait(self._h, C.byref(t)))
        if self.world == 1:
            return t.value // 2
        return self._comm.allreduce_sum(t.value) // 2

    def alive_mask(self):
        ns = int(self._scalar("nsp"))
        mask = np.zeros(ns, dtype=np.uint8)
        if ns:
            check(self._L.eco_alive_mask(self._h, _ptr(mask), ns))
        return mask

    def alive_mask_device(self):
        """The alive mask left on the device as a uint8 torch tensor aliasing the engine's buffer, or None when the kernels
        do not run on torch's current stream (the caller then uses alive_mask)."""
        import torch
        ns = int(self._scalar("nsp"))
        if not ns or not self._on_torch_stream():
            return None
        p = C.c_void_p()
        check(self._L.eco_alive_mask_device(self._h, C.byref(p), ns))
        return torch.as_tensor(_DevMem(p.value, ns, "|u1"), device=torch.device("cuda", self._device))

    def condense_apply(self, mask):
        r = C.c_uint()
        check(self._L.eco_condense_apply(self._h, _ptr(mask), len(mask), C.byref(r)))
        return r.value

    def mutate_cells(self):
        n = C.c_size_t()
        check(self._L.eco_mutate_cells(self._h, C.byref(n)))
        rec = (MutRecord * max(n.value, 1))()
        if n.value:
            check(self._L.eco_mutate_records(self._h, rec, n.value))
        return np.frombuffer(rec, dtype=np.uint32).reshape(-1, 3)[: n.value].copy()

    def mutate_model(self, parents):
        parents = np.ascontiguousarray(parents, dtype=np.uint32)
        check(self._L.eco_mutate_model(self._h, _ptr(parents), len(parents)))

    def lengths(self):
        d = {n: self._array_len(n, "f64") for n in self._TAILS_F64}
        d.update({n: self._array_len(n, "u32") for n in self._TAILS_U32})
        d["species"] = self._array_len("species", "i32")
        return d

    def tails(self, before):
        t = {k: self._get_array(k, "f64")[before[k]:] for k in self._TAILS_F64}
        t.update({k: self._get_array(k, "u32")[before[k]:] for k in self._TAILS_U32})
        t["species"] = self._get_array("species", "i32")[before["species"]:]
        return t

    def append_tails(self, tails):
        put = EcolabModelBase._set_array    # not an assignment by the script: the replicas stay marked identical
        for k in self._TAILS_F64:
            put(self, k, "f64", np.concatenate([self._get_array(k, "f64"), tails[k]]))
        for k in self._TAILS_U32:
            put(self, k, "u32", np.concatenate([self._get_array(k, "u32"), tails[k]]))
        put(self, "species", "i32", np.concatenate([self._get_array("species", "i32"), tails["species"]]))

    def append_species(self, cells):
        cells = np.ascontiguousarray(cells, dtype=np.uint32)
        check(self._L.eco_append_species(self._h, _ptr(cells), len(cells)))

    # ---- the reflected methods ---------------------------------------------------------------
    def migrate(self, wait=True):
        self._sync_replicas()
        if self.world == 1:
            t = C.c_uint64()
            check(self._L.eco_migrate(self._h, C.byref(t)))
            return t.value
        return dist_migrate(self, self._comm, wait)

    def step(self, niter=1, wait=True):
        """generate(niter) then migrate(); returns migrate's total.  wait=False enqueues both and returns None: no host
        round trip per step (read the total of the last step with last_migration())."""
        if self.world == 1:
            if not wait:
                check(self._L.eco_step_async(self._h, int(niter)))
                return None
            t = C.c_uint64()
            check(self._L.eco_step(self._h, int(niter), C.byref(t)))
            return t.value
        self.generate(niter)
        return self.migrate(wait)

    def step_host(self, dens_in, dens_out, niter=1):
        """generate(niter) + migrate() on a HOST-resident state: `dens_in` [owned cell][nsp] int32 is uploaded, the result
        lands in `dens_out`; copies and kernels are pipelined over bands of grid rows (eco_step_host).  Under
        torch.distributed each rank streams its own slab over its own PCIe link; the slab's two boundary rows are
        generated first so that the halo exchange does not wait for the upload."""
        a, b = np.ascontiguousarray(dens_in, dtype=np.int32), dens_out
        if b.dtype != np.int32 or not b.flags.c_contiguous or b.shape != a.shape:
            raise EcolabError("step_host: dens_out must be a C-contiguous int32 array shaped like dens_in")
        t = C.c_uint64()
        if self.world == 1:
            check(self._L.eco_step_host(self._h, _ptr(a), _ptr(b), a.shape[0], a.shape[1], int(niter), C.byref(t)))
            return t.value
        self._sync_replicas()
        check(self._L.eco_step_host_begin(self._h, _ptr(a), a.shape[0], a.shape[1], int(niter)))
        ht = self.halo_tensors()
        if ht is not None:
            self._comm.sendrecv_rows(*ht)
            self.halo_done()
        check(self._L.eco_step_host(self._h, _ptr(a), _ptr(b), a.shape[0], a.shape[1], int(niter), C.byref(t)))
        return self._comm.allreduce_sum(t.value) // 2

    def condense(self):
        if self.world == 1:
            return super().condense()
        self._sync_replicas()
        return dist_condense(self, self._comm)

    def mutate(self):
        if self.world == 1:
            return super().mutate()
        self._sync_replicas()
        dist_mutate(self, self._comm)
