"""GPU: seeded randomised differential test -- random grid shapes, species counts, row lengths, rates and random
SEQUENCES of the four stages (generate(niter) / migrate / condense / mutate) run through the C ABI and through the CPU
oracle side by side; every density, every return value and the species parameters must agree bit for bit after every
call.  The shapes are drawn to straddle the layout's corners: tile widths 32/16/8/4 (species capacity), ragged last
tiles (nx not a multiple of the tile width), one-row / one-column grids, fewer species-row blocks than warps, capacity
growth and re-layout under mutate, unit boundaries of migrate.  This stands in for compute-sanitizer, which the GPU pool
does not allow: an out-of-bounds or mis-indexed access in any of these corners changes a density somewhere."""
import numpy as np
import pytest

from helpers import SCALARS, densities, gpu_density, load_gpu, load_oracle, oracle_density, synth

pytestmark = pytest.mark.gpu


def draw_case(rng):
    nsp = int(rng.choice([1, 2, 7, 31, 33, 64, 100, 129, 300, 513, 600, 1000, 1030, 1800, 2100, 3600]))
    big = nsp > 700
    nx = int(rng.choice([1, 2, 3, 5, 15, 16, 17, 31, 33, 64, 70, 130])) if not big else int(rng.choice([1, 4, 9, 17, 33]))
    ny = int(rng.choice([1, 2, 3, 4, 7, 19, 40])) if not big else int(rng.choice([1, 2, 5]))
    k = int(rng.choice([0, 1, 2, 3, 4, 5, 8, 9, 13])) if nsp > 14 else int(rng.integers(0, max(1, nsp)))
    mig = float(rng.choice([0.0, 1e-3, 0.05, 0.11, 0.2499, 0.3, 2.0]))
    mut = float(rng.choice([0.0, 1e-3, 0.02]))
    dmax = int(rng.choice([2, 200, 200, 5000, 2 ** 20]))
    return (nx, ny), nsp, k, mig, mut, dmax


@pytest.mark.parametrize("case", range(48))
def test_random_op_sequences_match_the_oracle(oracle_mod, gpu, case):
    O = oracle_mod
    rng = np.random.default_rng(7000 + case)
    grid, nsp, k, mig, mut, dmax = draw_case(rng)
    p = synth(nsp, k, 100 + case, mig=mig, mut=mut)
    p["migration"] = p["migration"] * rng.uniform(0.0, 1.5, nsp)          # per-species rates, some beyond the 1/maxNbrs cap
    scal = dict(SCALARS, mut_max=max(mut, 1e-5))
    d = densities(grid[0] * grid[1], nsp, 100 + case, dmax=dmax)
    o = load_oracle(O, p, d, grid, seed=case, scalars=scal)
    m = load_gpu(gpu, p, d, grid, seed=case, scalars=scal)
    if rng.random() < 0.3:
        m.set_tuning("mig_rows", int(rng.choice([1, 2, 5, 16])))
    ops = list(rng.choice(["generate", "generate", "migrate", "migrate", "condense", "mutate"], size=7))
    ops[0] = "generate"                                                    # tstep > 0 before the first migrate / mutate
    for it, op in enumerate(ops):
        what = f"case {case}: grid {grid} nsp {nsp} nnz {k} step {it} {op}"
        if op == "generate":
            n = int(rng.choice([1, 1, 2, 5]))
            o.generate(n); m.generate(n)
        elif op == "migrate":
            assert m.migrate() == o.migrate(), what
        elif op == "condense":
            assert m.condense() == o.condense(), what
        else:
            O.srand(case * 10 + it); O.reset_grand(); o.mutate()
            gpu.lib().eco_array_urand_seed(case * 10 + it); gpu.lib().eco_array_grand_reset(); m.mutate()
            assert len(m.species) == o.nsp, what
            for name in ("repro_rate", "mutation", "migration"):
                assert np.array_equal(np.array(getattr(m, name)()), o.get(name)), what + " " + name
            for name in ("diag", "val", "row", "col"):
                assert np.array_equal(np.array(getattr(m.interaction, name)()), o.get("interaction." + name)), what + " " + name
        assert m.tstep() == int(o.get("tstep")), what
        if o.nsp:
            assert np.array_equal(gpu_density(m, grid), oracle_density(o, grid)), what
    m.close()
