"""Worker of tests/test_gpu_multi.py (one process per GPU under torchrun, NCCL): the row-slab partitioned
CUDA model against a single-process oracle run of the whole grid -- generate, halo migrate, condense with the
alive-mask all-reduce, mutate with the mutant all-gather -- gathered to rank 0 and compared bit for bit."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    import ecolab_b200 as E
    from helpers import SCALARS, densities, load_oracle, synth
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = E.Comm()
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)                 # kernels and NCCL on one stream: the configuration bench.py runs
    nx, ny, nsp, k, seed = 37, 19, int(os.environ.get("ECO_MULTI_NSP", "45")), 3, 7
    ndead = int(os.environ.get("ECO_MULTI_DEAD", "0"))
    p = synth(nsp, k, seed, mig=0.07, mut=2e-3)
    d = densities(nx * ny, nsp, seed)
    d[:, 4] = 0                                   # one species extinct everywhere -> condense removes it on every rank
    if ndead:                                     # ... and enough of them to cross a tile-width threshold: every rank re-tiles
        d[:, nsp - ndead:] = 0
    m = E.SpatialModel(local, comm)
    m.set_stream(stream)
    for kk, v in SCALARS.items():
        getattr(m, kk)(v)
    m.mut_max(2e-3)
    for kk, v in p.items():
        if kk.startswith("interaction."):
            getattr(m.interaction, kk.split(".")[1])(v)
        else:
            getattr(m, kk)(v)
    m.setGrid(nx, ny)
    m.seed(seed)
    m.set_density_array(d[m._y0 * nx:m._y1 * nx])

    def gathered():
        loc = torch.from_numpy(m.density_array()).cuda()
        ns = loc.shape[1]
        sizes = [(E.slab_bounds(ny, world, r)[1] - E.slab_bounds(ny, world, r)[0]) * nx for r in range(world)]
        outs = [torch.empty((s, ns), dtype=torch.int32, device="cuda") for s in sizes]
        dist.all_gather(outs, loc)
        return torch.cat(outs).cpu().numpy()

    o = None
    from oracle import oracle as O
    if rank == 0:
        O.build()
        sc = dict(SCALARS); sc["mut_max"] = 2e-3
        o = load_oracle(O, p, d, (nx, ny), seed=seed, scalars=sc)

    def check(what):
        g = gathered()
        if rank == 0:
            assert np.array_equal(g, o.get_density_all()), what

    for it in range(3):
        m.generate(1 + it)
        if rank == 0:
            o.generate(1 + it)
        check(f"generate {it}")
        moved = m.migrate()
        if rank == 0:
            assert moved == o.migrate(), f"migrate total {it}"
        check(f"migrate {it}")
    tw0 = int(m._scalar("tile_width"))
    removed = m.condense()
    if rank == 0:
        assert removed == o.condense() and removed >= 1 + ndead
    check("condense")
    if ndead:                                     # the layout changed under the peer-memory halo: it has to reconnect
        assert int(m._scalar("tile_width")) > tw0, "condense must have re-tiled"
        m.generate(1)
        if rank == 0:
            o.generate(1)
        moved = m.migrate()
        if rank == 0:
            assert moved == o.migrate(), "migrate total after the re-tiling"
        check("migrate after the re-tiling")
    # mutate: the cell-side speciation is deterministic (Philox); the parameter draws come from rank 0's libc stream
    # (ModelData::mutate runs on rank 0 only, :251-258), seeded here exactly like the oracle's -> everything must be equal
    before = int(gathered().sum())
    nsp0 = len(m.species)
    if rank == 0:
        O.srand(123); O.reset_grand()
        o.set("nprocs", world)   # species labels advance by nprocs per mutant (models/ecolab_model.cc:463-464)
        o.mutate()
        E.lib().eco_array_urand_seed(123); E.lib().eco_array_grand_reset()
    m.mutate()
    nsp1 = len(m.species)
    t = torch.tensor([nsp1], device="cuda")
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert int(lo) == int(hi) == nsp1 and nsp1 > nsp0, "every rank must append the same mutants"
    g = gathered()
    assert g.shape[1] == nsp1 and int(g.sum()) == before, "mutation moves individuals, it does not create them"
    if rank == 0:
        assert nsp1 == o.nsp
        assert np.array_equal(g, o.get_density_all()), "mutate densities"
    for name in ("repro_rate", "mutation", "migration", "create"):
        mine = torch.tensor(np.asarray(getattr(m, name)()), device="cuda")
        r0 = mine.clone(); dist.broadcast(r0, 0)
        assert torch.equal(mine, r0), f"{name} must be replicated identically"
        if rank == 0:
            assert np.array_equal(mine.cpu().numpy(), o.get(name)), f"{name} vs oracle"
    for name in ("diag", "val", "row", "col"):
        mine = torch.tensor(np.asarray(getattr(m.interaction, name)(), dtype=np.float64), device="cuda")
        r0 = mine.clone(); dist.broadcast(r0, 0)
        assert torch.equal(mine, r0), f"interaction.{name} must be replicated identically"
        if rank == 0:
            assert np.array_equal(mine.cpu().numpy(), o.get("interaction." + name).astype(np.float64)), f"interaction.{name} vs oracle"
    if rank == 0:
        assert m.species() == o.get("species").tolist()
    # the mutated model drives the next steps identically (capacity growth re-allocates the buffers: the peer-memory halo
    # has to reconnect), then condense again
    for it in range(2):
        m.generate(2)
        if rank == 0:
            o.generate(2)
        moved = m.migrate()
        if rank == 0:
            assert moved == o.migrate(), "migrate total after mutate"
        check(f"after mutate {it}")
    removed = m.condense()
    if rank == 0:
        assert removed == o.condense()
    check("second condense")
    # non-blocking steps (no host round trip per step), total fetched at the end
    for _ in range(3):
        m.step(1, wait=False)
        if rank == 0:
            o.generate(1); want_moved = o.migrate()
    moved = m.last_migration()
    if rank == 0:
        assert moved == want_moved
    check("async steps")
    # host-resident state streamed through every rank's own PCIe link (eco_step_host_begin / halo / eco_step_host)
    loc_in = m.density_array()
    loc_out = np.empty_like(loc_in)
    moved = m.step_host(loc_in, loc_out, 2)
    if rank == 0:
        o.generate(2)
        assert moved == o.migrate(), "step_host total"
    check("step_host device state")
    assert np.array_equal(loc_out, m.density_array()), "step_host host result"
    # the scripts' view of a distributed model: whole-grid nsp(), collective cell reads, dropped remote writes
    ns_all = m.nsp()
    if rank == 0:
        assert ns_all == o.nsp_per_cell().tolist()
    probe = m.cell(nx - 1, ny - 1).density()          # owned by the last rank, read everywhere
    if rank == 0:
        assert probe == o.get_density((nx - 1) + nx * (ny - 1)).tolist()
    m.cell(0, 0).density(probe)                       # every rank executes it; only the owner (rank 0) stores it
    if rank == 0:
        o.set_density(0, np.array(probe, dtype=np.int32))
    check("spmd cell write")
    halo = "peer" if int(m._scalar("peer_connected")) else "nccl"
    want_halo = os.environ.get("ECOLAB_B200_HALO", "peer")
    assert halo == want_halo or m._peer_failed, (halo, want_halo)
    dist.barrier()
    if rank == 0:
        print("MULTI_GPU_OK", world, halo)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
