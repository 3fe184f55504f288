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
    nx, ny, nsp, k, seed = 37, 11, 45, 3, 7
    p = synth(nsp, k, seed, mig=0.07, mut=2e-3)
    d = densities(nx * ny, nsp, seed)
    d[:, 4] = 0                                   # one species extinct everywhere -> condense removes it on every rank
    m = E.SpatialModel(local, comm)
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
    if rank == 0:
        from oracle import oracle as O
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
    removed = m.condense()
    if rank == 0:
        assert removed == o.condense() and removed >= 1
    check("condense")
    # mutate: the cell-side speciation is deterministic (Philox); the parameter draws use rank 0's libc stream,
    # so compare what must agree: species counts on every rank, conservation of individuals, parents' loss
    before = int(gathered().sum())
    nsp0 = len(m.species)
    m.mutate()
    nsp1 = len(m.species)
    t = torch.tensor([nsp1], device="cuda")
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert int(lo) == int(hi) == nsp1 and nsp1 > nsp0, "every rank must append the same mutants"
    g = gathered()
    assert g.shape[1] == nsp1 and int(g.sum()) == before, "mutation moves individuals, it does not create them"
    assert (g[:, nsp0:].sum(axis=0) == 1).all(), "each mutant is born as one individual in one cell"
    rr = torch.tensor(np.asarray(m.repro_rate()), device="cuda")
    r0 = rr.clone(); dist.broadcast(r0, 0)
    assert torch.equal(rr, r0), "species parameters must be replicated identically"
    m.generate(2); moved = m.migrate()
    g = gathered()
    assert (g >= 0).all()
    dist.barrier()
    if rank == 0:
        print("MULTI_GPU_OK", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
