"""Test-only: one rank of the distributed model emulated on the CPU with the oracle's slab halves,
driven by the PRODUCT's distributed algorithms (ecolab_b200.spatial.dist_*) over gloo.  The result,
gathered from all ranks, must equal a single-process oracle run of the whole grid."""
import numpy as np
import torch

from ecolab_b200.spatial import dist_condense, dist_migrate, dist_mutate, slab_bounds
from helpers import SCALARS, densities, load_oracle, synth

F64 = ("create", "repro_rate", "mutation", "migration", "interaction.diag", "interaction.val")
U32 = ("interaction.row", "interaction.col")


class SlabRank:
    """The `ops` protocol of ecolab_b200.spatial over an oracle slab."""
    _TAILS_F64, _TAILS_U32 = F64, U32     # the arrays ModelData::mutate appends to (dist_mutate broadcasts their tails)

    def __init__(self, O, p, dens, nx, ny, comm, seed):
        self.O, self.nx, self.ny, self.comm = O, nx, ny, comm
        self.y0, self.y1 = slab_bounds(ny, comm.world, comm.rank)
        o = O.Oracle(rng=O.RNG_PHILOX, seed=seed)
        for k, v in SCALARS.items():
            o.set(k, v)
        for k, v in p.items():
            o.set(k, v)
        o.set("nprocs", comm.world)
        o.set_grid_slab(nx, ny, self.y0, self.y1 - self.y0)
        o.set_density_all(dens[self.y0 * nx:self.y1 * nx])
        self.o = o

    # halo: CPU tensors stand in for the device rows
    def halo_tensors(self):
        d = self.o.get_density_all()
        ns = d.shape[1]
        self._top = torch.zeros(self.nx * ns, dtype=torch.int32)
        self._bot = torch.zeros(self.nx * ns, dtype=torch.int32)
        first = torch.from_numpy(d[: self.nx].copy().ravel())
        last = torch.from_numpy(d[-self.nx:].copy().ravel())
        return first, self._top, last, self._bot

    def halo_done(self):
        self.o.set_ghost(0, self._top.numpy())
        self.o.set_ghost(1, self._bot.numpy())

    def migrate_local(self):
        return self.o.migrate_local()

    def alive_mask(self):
        return self.o.alive_mask()

    def condense_apply(self, mask):
        return self.o.condense_apply(mask)

    def mutate_cells(self):
        return self.o.mutate_cells()

    def mutate_model(self, parents):
        self.o.mutate_model(parents)

    def lengths(self):
        d = {k: len(self.o.get(k)) for k in F64 + U32}
        d["species"] = len(self.o.get("species"))
        return d

    def tails(self, before):
        return {k: self.o.get(k)[before[k]:] for k in F64 + U32 + ("species",)}

    def append_tails(self, tails):
        for k in F64 + U32 + ("species",):
            self.o.set(k, np.concatenate([self.o.get(k), tails[k]]))

    def append_species(self, cells):
        self.o.append_species(cells)


def run_distributed_against_oracle(comm, nx=5, ny=7, nsp=19, seed=11):
    from oracle import oracle as O
    p = synth(nsp, 3, seed, mut=0.02)
    scal = dict(SCALARS, mut_max=0.02)
    d = densities(nx * ny, nsp, seed)
    d[:, 4] = 0
    r = SlabRank(O, p, d, nx, ny, comm, seed)
    r.o.set("mut_max", 0.02)
    full = load_oracle(O, p, d, (nx, ny), seed=seed, scalars=scal)
    full.set("nprocs", comm.world)   # species labels advance by nprocs per mutant (models/ecolab_model.cc:464-465)

    def gathered():
        mine = torch.from_numpy(r.o.get_density_all().copy())
        sizes = [(slab_bounds(ny, comm.world, k)[1] - slab_bounds(ny, comm.world, k)[0]) * nx for k in range(comm.world)]
        outs = [torch.zeros((s, mine.shape[1]), dtype=torch.int32) for s in sizes]
        comm.dist.all_gather(outs, mine) if len(set(sizes)) == 1 else None
        if len(set(sizes)) != 1:
            objs = [None] * comm.world
            comm.dist.all_gather_object(objs, mine.numpy())
            return np.concatenate(objs)
        return torch.cat(outs).numpy()

    for it in range(3):
        r.o.generate(2); full.generate(2)
        assert np.array_equal(gathered(), full.get_density_all()), "generate"
        moved = dist_migrate(r, comm)
        assert moved == full.migrate(), "migrate total"
        assert np.array_equal(gathered(), full.get_density_all()), "migrate"
        O.srand(77 + it); O.reset_grand()
        n_new = dist_mutate(r, comm)
        O.srand(77 + it); O.reset_grand()
        full.mutate()
        assert r.o.nsp == full.nsp and (it > 0 or n_new > 0), "mutate count"
        assert np.array_equal(gathered(), full.get_density_all()), "mutate densities"
        for k in F64 + U32 + ("species",):
            assert np.array_equal(r.o.get(k), full.get(k)), ("mutate " + k)
        assert dist_condense(r, comm) == full.condense(), "condense count"
        assert np.array_equal(gathered(), full.get_density_all()), "condense"
