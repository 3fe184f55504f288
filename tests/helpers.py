"""Shared builders for the parity tests: one seeded synthetic model, loaded
identically into the oracle and into the CUDA engine."""
import numpy as np


def synth(nsp, nnz_per_row, seed, density="const", dmax=200, mig=0.1, mut=1e-5, dense=False):
    """Species parameters as the reference scripts choose them (models/spatial_ecolab.py:17-56):
    repro_rate ~ U(-0.1,0.1), diag = -1e-3, val ~ U(-1e-5,1e-5), k distinct off-diagonal columns per row
    (entries deliberately NOT sorted by row, to exercise the COO-order rule of computeODiagIdx)."""
    rng = np.random.default_rng(seed)
    p = {}
    p["species"] = np.arange(nsp, dtype=np.int32)
    p["repro_rate"] = rng.uniform(-0.1, 0.1, nsp)
    p["interaction.diag"] = np.full(nsp, -1e-3)
    rows, cols = [], []
    for i in range(nsp):
        k = min(nnz_per_row, nsp - 1)
        if dense:
            c = np.delete(np.arange(nsp), i)
        else:
            c = rng.choice(nsp - 1, size=k, replace=False) if k else np.zeros(0, dtype=np.int64)
            c = c + (c >= i)
        rows.append(np.full(len(c), i))
        cols.append(c)
    row = np.concatenate(rows).astype(np.uint32) if rows else np.zeros(0, np.uint32)
    col = np.concatenate(cols).astype(np.uint32) if cols else np.zeros(0, np.uint32)
    perm = rng.permutation(len(row))
    p["interaction.row"], p["interaction.col"] = row[perm], col[perm]
    p["interaction.val"] = rng.uniform(-1e-5, 1e-5, len(row))
    p["mutation"] = np.full(nsp, mut)
    p["migration"] = np.full(nsp, mig)
    p["create"] = np.zeros(nsp)
    return p


def densities(ncells, nsp, seed, kind="rand", dmax=200):
    rng = np.random.default_rng(seed + 1000)
    if kind == "const":
        return np.full((ncells, nsp), 100, dtype=np.int32)
    d = rng.integers(0, dmax, size=(ncells, nsp), dtype=np.int32)
    d[rng.random((ncells, nsp)) < 0.1] = 0          # exercise ROUND's x<=0 branch
    return d


SCALARS = dict(sp_sep=0.1, mut_max=1e-5, repro_min=-0.1, repro_max=0.1, odiag_min=-1e-5, odiag_max=1e-5, gen_bias=0.0)


def load_oracle(O, p, dens, grid=None, seed=10, scalars=SCALARS):
    o = O.Oracle(rng=O.RNG_PHILOX, seed=seed)
    for k, v in scalars.items():
        o.set(k, v)
    for k, v in p.items():
        o.set(k, v)
    if grid:
        o.set_grid(*grid)
        o.set_density_all(dens)
    else:
        o.set_density(0, dens[0])
    return o


def load_gpu(E, p, dens, grid=None, seed=10, scalars=SCALARS, device=0):
    m = E.SpatialModel(device) if grid else E.PanmicticModel(device)
    for k, v in scalars.items():
        getattr(m, k)(v)
    for k, v in p.items():
        if k.startswith("interaction."):
            getattr(m.interaction, k.split(".")[1])(v)
        else:
            getattr(m, k)(v)
    if grid:
        m.setGrid(*grid)
        m.set_density_array(dens)
    else:
        m.density(dens[0])
    m.seed(seed)
    return m


def gpu_density(m, grid):
    if grid:
        return m.density_array()
    return np.array([m.density()], dtype=np.int32)


def oracle_density(o, grid):
    if grid:
        return o.get_density_all()
    return np.array([o.get_density(0)], dtype=np.int32)


# ---------------------------------------------------------------------------------------------
# band parity at sizes the oracle cannot hold whole (BASELINE.json C3 / C4): generate is independent per cell and
# migrate reaches one row up and down, so rows [y, y+nb) after generate(1)+migrate() depend only on rows [y-1, y+nb+1)
# before it.  The oracle runs that 6-row slab with the true global row numbers (the Philox stream is keyed on them).
# ---------------------------------------------------------------------------------------------
def collect_rows(m, y, n):
    """Grid rows [y, y+n) as [n*nx][nsp] int32 on rank 0 (None elsewhere); the rows may straddle rank seams."""
    world = getattr(m, "world", 1)
    if world == 1:
        return m.density_rows(y, n)
    import torch.distributed as dist
    a, b = max(y, m._y0), min(y + n, m._y1)
    part = (a, m.density_rows(a, b - a)) if b > a else None
    parts = [None] * world if m.rank == 0 else None
    dist.gather_object(part, parts, dst=0)
    if m.rank != 0:
        return None
    parts = sorted((q for q in parts if q is not None), key=lambda q: q[0])
    return np.concatenate([q[1] for q in parts])


def band_step_equals_oracle(O, m, p, scalars, seed, y, nb=4):
    """Run ONE generate(1)+migrate() step on the GPU model `m` (any size, any partition) and check rows [y, y+nb) of the
    result against the oracle fed with the GPU's own pre-step rows [y-1, y+nb+1).  Collective under torch.distributed;
    returns (equal, moved) on rank 0 and (None, moved) elsewhere."""
    nx, ny = m._nx, m._ny
    ya, yb = max(0, y - 1), min(ny, y + nb + 1)
    pre = collect_rows(m, ya, yb - ya)
    tstep, last_mig = m.tstep(), m.last_mig_tstep()
    moved = m.step(1)
    post = collect_rows(m, y, nb)
    if pre is None:
        return None, moved
    ns = pre.shape[1]
    o = O.Oracle(rng=O.RNG_PHILOX, seed=seed)
    for k, v in scalars.items():
        o.set(k, v)
    for k, v in p.items():
        o.set(k, v)
    o.set_grid_slab(nx, ny, ya, yb - ya)
    o.set_density_all(pre)
    zeros = np.zeros((nx, ns), dtype=np.int32)
    o.set_ghost(0, zeros); o.set_ghost(1, zeros)      # only reach the slab's outer rows, which are not compared
    o.make_consistent()
    o.set("tstep", tstep); o.set("last_mig_tstep", last_mig)
    o.generate(1)
    o.migrate_local()
    want = o.get_density_all()[(y - ya) * nx:(y - ya + nb) * nx]
    return bool(np.array_equal(post, want)), moved
