"""GPU: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): integer densities bit-exact for generate + migrate + condense under the
shared Philox stream; the pre-rounding float update within 1e-6 relative (here: exactly equal, the kernel
keeps the reference's FP64 operation order); mutation statistically equivalent (and, because product and
oracle draw from the same libc rand() stream inside this process, also bit-equal when re-seeded alike).
"""
import os

import numpy as np
import pytest

from helpers import SCALARS, densities, gpu_density, load_gpu, load_oracle, oracle_density, synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

# (grid, nsp, nnz/row): ragged shapes on purpose -- nx not a multiple of the 32-cell tile, nsp not a multiple of 4
SHAPES = [
    (None, 2, 1), (None, 100, 3), (None, 1, 0), (None, 1023, 5),
    ((1, 1), 7, 2), ((2, 2), 30, 3), ((8, 5), 37, 3), ((33, 3), 64, 16), ((64, 2), 9, 8), ((70, 9), 129, 3),
    ((31, 1), 5, 0), ((1, 40), 12, 4),
    # long rows: the software-pipelined term loop (W >= 8) with every remainder, stream in shared memory and not
    (None, 50, 8), (None, 50, 9), ((5, 5), 80, 11), ((9, 2), 64, 12), ((20, 3), 200, 51), ((18, 2), 1100, 23),
    # tile widths above 1024 species: 16 cells (<= 1784), 8 cells (<= 3568), 4 cells beyond
    ((19, 2), 1700, 4), ((10, 3), 2100, 7), ((5, 2), 3700, 2),
]


def pair(oracle_mod, gpu, grid, nsp, k, seed=5, **kw):
    ncells = grid[0] * grid[1] if grid else 1
    p = synth(nsp, k, seed, **kw)
    d = densities(ncells, nsp, seed)
    return load_oracle(oracle_mod, p, d, grid, seed=seed), load_gpu(gpu, p, d, grid, seed=seed), grid


@pytest.mark.parametrize("grid,nsp,k", SHAPES)
def test_generate_bit_exact(oracle_mod, gpu, grid, nsp, k):
    o, m, grid = pair(oracle_mod, gpu, grid, nsp, k)
    for niter in (1, 1, 3, 10):        # niter==1: single-pass kernel; niter>1: fused multi-step kernel
        o.generate(niter); m.generate(niter)
        assert np.array_equal(gpu_density(m, grid), oracle_density(o, grid)), f"niter={niter}"
        assert m.tstep() == int(o.get("tstep"))


@pytest.mark.parametrize("grid,nsp,k", [((8, 5), 37, 3), ((33, 3), 64, 5), ((70, 9), 129, 3), ((17, 4), 1000, 3), ((5, 3), 8, 1),
                                        ((9, 2), 200, 0)])
@pytest.mark.parametrize("knob,val", [("gen_dual", -1), ("gen_nobsmem", 1), ("gen_stages", 1)])
def test_generate_kernel_forms(oracle_mod, gpu, grid, nsp, k, knob, val):
    """generate(1) through each shipped form of the pipeline kernel (one species-row block per consumer warp instead of
    two, beta through L1/L2 instead of shared memory, a single tile buffer): same bits, odd block counts included."""
    o, m, grid = pair(oracle_mod, gpu, grid, nsp, k)
    m.set_tuning(knob, val)
    for niter in (1, 1, 2):
        o.generate(niter); m.generate(niter)
        assert np.array_equal(gpu_density(m, grid), oracle_density(o, grid))


def test_generate_beta_values_outside_the_exact_fma_range(oracle_mod, gpu):
    """The product val*n is one fused multiply-add against -(val*2^52) (k_generate.cuh); that needs val = 0 or a normal
    number.  Subnormal and huge entries switch the engine to the multiply-then-add kernel: same bits either way."""
    grid, nsp = (9, 4), 40
    p = synth(nsp, 4, 13)
    p["interaction.val"][::7] = 5e-324          # smallest subnormal
    p["interaction.val"][3::11] = -2.5e-310
    p["interaction.val"][5::13] = 0.0
    d = densities(36, nsp, 13)
    o = load_oracle(oracle_mod, p, d, grid); m = load_gpu(gpu, p, d, grid)
    for niter in (1, 3):
        o.generate(niter); m.generate(niter)
        assert np.array_equal(m.density_array(), o.get_density_all())


@pytest.mark.parametrize("grid,nsp,k", [(None, 100, 3), ((8, 5), 37, 3), ((33, 3), 64, 16)])
def test_pre_rounding_float_update(oracle_mod, gpu, grid, nsp, k):
    o, m, grid = pair(oracle_mod, gpu, grid, nsp, k)
    o.generate(2); m.generate(2)
    x_gpu = m.pre_rounding()
    o.set("record_pre", 1); o.generate(1)
    x_cpu = o.pre_rounding().reshape(x_gpu.shape)
    rel = np.abs(x_gpu - x_cpu) / np.maximum(np.abs(x_cpu), 1e-300)
    assert rel.max() <= 1e-6            # the stated tolerance
    assert np.array_equal(x_gpu, x_cpu)  # and in fact identical: same operation order, no FMA


def test_generate_round_extremes(oracle_mod, gpu):
    """ROUND's three branches (models/ecolab_model.cc:57-66): x<=0 -> 0, x>=INT_MAX-1 -> INT_MAX-1, else stochastic."""
    nsp = 8
    p = synth(nsp, 2, 9)
    p["repro_rate"] = np.array([5.0, -3.0, 0.5, -1.0, 0.0, 1e-9, 0.999, -0.999])
    p["interaction.diag"] = np.zeros(nsp); p["interaction.val"][:] = 0
    d = np.array([[2 ** 31 - 2, 50, 2 ** 30, 17, 0, 1, 2 ** 31 - 2, 3]] * 40, dtype=np.int32)
    o = load_oracle(oracle_mod, p, d, (40, 1), seed=3); m = load_gpu(gpu, p, d, (40, 1), seed=3)
    for _ in range(3):
        o.generate(1); m.generate(1)
        assert np.array_equal(m.density_array(), o.get_density_all())
    got = m.density_array()
    assert (got[:, 0] == 2 ** 31 - 2).all() and (got[:, 1] == 0).all() and (got[:, 3] == 0).all()


@pytest.mark.parametrize("grid,nsp,k", [s for s in SHAPES if s[0]])
def test_migrate_bit_exact_and_conserves(oracle_mod, gpu, grid, nsp, k):
    o, m, grid = pair(oracle_mod, gpu, grid, nsp, k, mig=0.07)
    for it in range(3):
        o.generate(1 + it); m.generate(1 + it)
        before = m.species_totals()
        mv_o, mv_g = o.migrate(), m.migrate()
        assert mv_g == mv_o
        dg = m.density_array()
        assert np.array_equal(dg, o.get_density_all())
        assert np.array_equal(m.species_totals(), before) and (dg >= 0).all()   # test/00/spatial_ecolab.sh:56-77
        assert m.last_mig_tstep() == m.tstep()


@pytest.mark.parametrize("grid,nsp,k", [((250, 3), 1000, 3), ((256, 4), 37, 3), ((120, 5), 300, 2), ((64, 7), 2000, 1), ((40, 70), 1000, 3),
                                        ((33, 150), 64, 2), ((7, 41), 2100, 1), ((6, 20), 3700, 1)])
@pytest.mark.parametrize("rows", [0, 1, 7, 100000])
def test_migrate_kernels(oracle_mod, gpu, grid, nsp, k, rows):
    """migrate with the automatic unit length and with forced ones (1 row: every row is a unit boundary; 7; the whole
    slab): same bits, ragged last tiles, partial species groups (fewer species-row blocks than warps in a CTA) and tile
    widths 32/16/8/4 included.  Units are handed out dynamically, so the order they run in differs from launch to launch."""
    o, m, grid = pair(oracle_mod, gpu, grid, nsp, k, mig=0.11)
    if rows:
        m.set_tuning("mig_rows", rows)
    for it in range(2):
        o.generate(1); m.generate(1)
        assert m.migrate() == o.migrate()
        assert np.array_equal(m.density_array(), o.get_density_all())


def test_migrate_extreme_densities(oracle_mod, gpu):
    """Densities at the top of the int32 range next to empty cells: the 2^52+n trick and the flux sums stay exact, totals
    are conserved, nothing goes negative (the cap 1/maxNbrs guarantees a cell is never overdrawn, :626-629)."""
    grid, nsp = (37, 5), 9
    p = synth(nsp, 2, 6)
    p["migration"] = np.array([1.0, 0.25, 0.2499999, 0.1, 1e-9, 0.5, 0.05, 0.24, 3.0])
    rng = np.random.default_rng(11)
    d = np.zeros((grid[0] * grid[1], nsp), dtype=np.int32)
    hot = rng.random(d.shape) < 0.3
    d[hot] = rng.integers(2 ** 31 - 1000, 2 ** 31 - 2, size=int(hot.sum()), dtype=np.int64).astype(np.int32)
    d[rng.random(d.shape) < 0.1] = 2 ** 31 - 2
    o = load_oracle(oracle_mod, p, d, grid); m = load_gpu(gpu, p, d, grid)
    o.set("tstep", 1); m.tstep(1)            # dt = 1 since last_mig_tstep = 0
    for rows in (0, 3, 0):
        m.set_tuning("mig_rows", rows)
        before = m.species_totals()
        assert m.migrate() == o.migrate()
        got = m.density_array()
        assert np.array_equal(got, o.get_density_all()) and (got >= 0).all()
        assert np.array_equal(m.species_totals(), before)
        o.set("tstep", int(o.get("tstep")) + 1); m.tstep(m.tstep() + 1)


def test_migrate_cap_and_per_species_rates(oracle_mod, gpu):
    """cm = min(dt*migration, 1/maxNbrs) (:626-629): huge rates are capped so a cell is never overdrawn."""
    grid, nsp = (9, 6), 11
    p = synth(nsp, 2, 4)
    p["migration"] = np.array([0, 1e-3, 0.1, 0.24, 0.25, 0.26, 1.0, 50.0, 1e-9, 0.2, 0.05])
    d = densities(54, nsp, 4, dmax=100000)
    o = load_oracle(oracle_mod, p, d, grid); m = load_gpu(gpu, p, d, grid)
    for _ in range(4):
        o.generate(3); m.generate(3)
        assert m.migrate() == o.migrate()
        assert np.array_equal(m.density_array(), o.get_density_all())
    assert (m.density_array() >= 0).all()


def test_step_equals_generate_then_migrate(oracle_mod, gpu):
    o, m, grid = pair(oracle_mod, gpu, (40, 7), 50, 3, mig=0.1)
    for niter in (1, 4):
        o.generate(niter); mv = o.migrate()
        assert m.step(niter) == mv
        assert np.array_equal(m.density_array(), o.get_density_all())


@pytest.mark.parametrize("grid,nsp,k,niter", [((70, 9), 129, 3, 1), ((33, 40), 64, 16, 1), ((8, 5), 37, 3, 4), ((5, 1), 9, 2, 1)])
def test_step_host_banded_pipeline(oracle_mod, gpu, grid, nsp, k, niter):
    """eco_step_host = set_density_all + generate(niter) + migrate + get_density_all, streamed in row bands."""
    o, m, grid = pair(oracle_mod, gpu, grid, nsp, k, mig=0.07)
    d_in = oracle_density(o, grid).copy()
    out = np.empty_like(d_in)
    for it in range(3):
        o.set_density_all(d_in)
        o.generate(niter); want_moved = o.migrate()
        moved = m.step_host(d_in, out, niter)
        assert moved == want_moved
        assert np.array_equal(out, o.get_density_all()), f"step {it}"
        assert np.array_equal(m.density_array(), out)            # the device state is the same result
        assert m.tstep() == int(o.get("tstep")) and m.last_mig_tstep() == m.tstep()
        d_in = out.copy()


@pytest.mark.parametrize("grid,nsp,k", [((70, 300), 129, 3), ((1030, 64), 300, 2), ((5, 1), 9, 2)])
def test_step_host_in_place(oracle_mod, gpu, grid, nsp, k):
    """One host buffer as input AND output of eco_step_host (what bench.py's e2e leg does at the metric's grid, where a
    second 64 GiB copy would not fit the box): band k-1 is written back only after band k has been read.  Many bands
    (the grid is tall enough for the 8-band minimum to bite), and a one-row grid."""
    o, m, grid = pair(oracle_mod, gpu, grid, nsp, k, mig=0.07)
    buf = oracle_density(o, grid).copy()
    for it in range(3):
        o.generate(1); want_moved = o.migrate()
        assert m.step_host(buf, buf, 1) == want_moved
        assert np.array_equal(buf, o.get_density_all()), f"step {it}"
        assert np.array_equal(m.density_array(), buf)


@pytest.mark.parametrize("grid,nsp,k", [(None, 100, 3), ((8, 5), 37, 3), ((70, 9), 129, 3)])
def test_condense_bit_exact(oracle_mod, gpu, grid, nsp, k):
    ncells = grid[0] * grid[1] if grid else 1
    p = synth(nsp, k, 6)
    d = densities(ncells, nsp, 6)
    d[:, [0, nsp // 2, nsp - 1]] = 0
    d[:, 3] = 0
    o = load_oracle(oracle_mod, p, d, grid); m = load_gpu(gpu, p, d, grid)
    o.generate(2); m.generate(2)
    removed = o.condense()
    assert m.condense() == removed and removed >= 4
    assert np.array_equal(gpu_density(m, grid), oracle_density(o, grid))
    for name in ("repro_rate", "mutation", "migration", "create"):
        assert np.array_equal(np.array(getattr(m, name)()), o.get(name))
    for name in ("diag", "val", "row", "col"):
        assert np.array_equal(np.array(getattr(m.interaction, name)()), o.get("interaction." + name)), name
    assert m.species() == o.get("species").tolist()
    assert m.condense() == o.condense() == 0
    o.generate(5); m.generate(5)           # the compacted beta drives the next steps identically
    assert np.array_equal(gpu_density(m, grid), oracle_density(o, grid))


@pytest.mark.parametrize("grid,nsp,ndead,tw_before,tw_after", [((24, 5), 600, 150, 16, 32), ((20, 3), 1900, 300, 8, 16),
                                                               ((19, 2), 3650, 200, 4, 8)])
def test_condense_retiles_to_the_wider_tile(oracle_mod, gpu, grid, nsp, ndead, tw_before, tw_after):
    """The blocked layout follows the species count both ways: condense() below a tile-width threshold re-lays the
    densities out on the wider tile, the next mutate() that outgrows the capacity goes back -- same bits throughout."""
    ncells = grid[0] * grid[1]
    O = oracle_mod
    p = synth(nsp, 2, 12, mig=0.05, mut=400.0 / (ncells * nsp))     # ~100 mutants per mutate() and unit of time
    d = densities(ncells, nsp, 12)
    d[:, nsp - ndead:] = 0
    d[:, 7] = 0
    o = load_oracle(oracle_mod, p, d, grid, seed=12); m = load_gpu(gpu, p, d, grid, seed=12)
    try:
        assert int(m._scalar("tile_width")) == tw_before
        o.generate(1); m.generate(1)
        removed = o.condense()
        assert m.condense() == removed and removed >= ndead
        assert int(m._scalar("tile_width")) == tw_after and int(m._scalar("capacity")) >= nsp - removed
        assert np.array_equal(m.density_array(), o.get_density_all())
        for it in range(3):
            o.generate(2); m.generate(2)
            assert o.migrate() == m.migrate()
            assert np.array_equal(m.density_array(), o.get_density_all())
            assert np.array_equal(m.species_totals(), o.get_density_all().astype(np.int64).sum(axis=0))
        nb = int(m._scalar("nsp"))
        for it in range(4):                       # speciation until the capacity (and with it the tile width) changes again
            o.generate(1); m.generate(1)
            O.srand(300 + it); O.reset_grand(); o.mutate()
            gpu.lib().eco_array_urand_seed(300 + it); gpu.lib().eco_array_grand_reset(); m.mutate()
            assert int(m._scalar("nsp")) == o.nsp
            assert np.array_equal(m.density_array(), o.get_density_all())
        assert o.nsp > nb and int(m._scalar("tile_width")) <= tw_after and int(m._scalar("capacity")) >= o.nsp
        o.generate(2); m.generate(2)
        assert o.migrate() == m.migrate()
        assert np.array_equal(m.density_array(), o.get_density_all())
    finally:
        m.close()


def test_condense_sh(gpu):
    """test/00/condense.sh:14-21 through the reference-facing API."""
    m = gpu.PanmicticModel(0)
    m.density([1, 0, 1, 0, 1])
    m.makeConsistent()
    m.repro_rate([1, 1, 1, 1, 1])
    m.interaction.diag([1, 1, 1, 1, 1])
    m.condense()
    assert m.density() == [1, 1, 1]


@pytest.mark.parametrize("grid,nsp", [(None, 60), ((8, 5), 37), ((70, 9), 40)])
def test_mutate_same_streams_bit_exact(oracle_mod, gpu, grid, nsp):
    """Cell side: same Philox stream -> same speciation counts.  Species side (ModelData::mutate): product and oracle
    both consume libc rand(); re-seeding before each makes the parameter draws identical too."""
    O = oracle_mod
    ncells = grid[0] * grid[1] if grid else 1
    p = synth(nsp, 3, 8, mut=0.02)
    scal = dict(SCALARS, mut_max=0.02)
    d = densities(ncells, nsp, 8)
    o = load_oracle(O, p, d, grid, scalars=scal); m = load_gpu(gpu, p, d, grid, scalars=scal)
    for it in range(3):
        o.generate(3); m.generate(3)
        O.srand(100 + it); O.reset_grand(); o.mutate()
        gpu.lib().eco_array_urand_seed(100 + it); gpu.lib().eco_array_grand_reset(); m.mutate()
        assert len(m.species) == o.nsp and o.nsp > nsp
        assert np.array_equal(gpu_density(m, grid), oracle_density(o, grid))
        for name in ("repro_rate", "mutation", "migration", "create"):
            assert np.array_equal(np.array(getattr(m, name)()), o.get(name)), name
        for name in ("diag", "val", "row", "col"):
            assert np.array_equal(np.array(getattr(m.interaction, name)()), o.get("interaction." + name)), name
        assert m.species() == o.get("species").tolist()
        assert m.last_mut_tstep() == m.tstep()
    o.generate(2); m.generate(2)
    assert np.array_equal(gpu_density(m, grid), oracle_density(o, grid))


def test_mutate_distribution_ks(oracle_mod, gpu):
    """Statistical equivalence with DIFFERENT random streams (north_star: KS on species counts and traits).
    array_grand is a running-CLT generator (src/random.cc:59-64) whose successive values are strongly correlated, so
    samples are only i.i.d. ACROSS trials: every trial restarts it, and one trait per trial enters the test."""
    from scipy.stats import ks_2samp
    O = oracle_mod
    nsp, grid = 50, (16, 16)
    p = synth(nsp, 3, 12, mut=0.01)
    scal = dict(SCALARS, mut_max=0.01)
    d = densities(256, nsp, 12)
    counts_o, counts_g, r_o, r_g, mu_o, mu_g = [], [], [], [], [], []
    for trial in range(60):
        o = load_oracle(O, p, d, grid, seed=1000 + trial, scalars=scal)
        m = load_gpu(gpu, p, d, grid, seed=5000 + trial, scalars=scal)
        o.set("tstep", 10); m.tstep(10)
        O.srand(trial); O.reset_grand(); o.mutate()
        gpu.lib().eco_array_urand_seed(7777 + trial); gpu.lib().eco_array_grand_reset(); m.mutate()
        counts_o.append(o.nsp - nsp); counts_g.append(len(m.species) - nsp)
        r_o.append(o.get("repro_rate")[-1]); r_g.append(m.repro_rate()[-1])
        mu_o.append(o.get("mutation")[-1]); mu_g.append(m.mutation()[-1])
    assert min(counts_o) > 0 and min(counts_g) > 0
    assert ks_2samp(counts_o, counts_g).pvalue > 0.01
    assert ks_2samp(r_o, r_g).pvalue > 0.01
    assert ks_2samp(mu_o, mu_g).pvalue > 0.01


def test_fill_density_random_matches_oracle(oracle_mod, gpu):
    o, m, grid = pair(oracle_mod, gpu, (70, 9), 129, 3)
    o.fill_density_random(3, 250); m.fill_density_random(3, 250)
    d = m.density_array()
    assert np.array_equal(d, o.get_density_all()) and d.min() >= 3 and d.max() <= 250 and d.std() > 50
    o.generate(2); m.generate(2)
    assert m.migrate() == o.migrate() > 0
    assert np.array_equal(m.density_array(), o.get_density_all())


def test_golden_fixtures(gpu):
    """The committed fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py from the oracle)."""
    from golden.make_golden import CASES, case_inputs, run_script
    for name in CASES:
        grid, p, d, script = case_inputs(name)
        m = load_gpu(gpu, p, d, grid, seed=10 + len(name))
        ops = {"generate": m.generate, "migrate": getattr(m, "migrate", None), "condense": m.condense,
               "density": lambda: gpu_density(m, grid)}
        got = run_script(ops, script)
        want = np.load(os.path.join(GOLDEN, name + ".npz"))
        for k in want.files:
            assert np.array_equal(want[k], got[k]), (name, k)


def test_script_cadence_matches_oracle(oracle_mod, gpu):
    """models/spatial_ecolab.py:66-78 cadence: generate(100); mutate(); migrate(); condense() -- several rounds."""
    O = oracle_mod
    grid, nsp = (12, 12), 100
    p = synth(nsp, 3, 21, mut=1e-4)
    scal = dict(SCALARS, mut_max=1e-4)
    d = densities(144, nsp, 21, kind="const")
    o = load_oracle(O, p, d, grid, scalars=scal); m = load_gpu(gpu, p, d, grid, scalars=scal)
    for it in range(4):
        o.generate(100); m.generate(100)
        O.srand(it); O.reset_grand(); o.mutate()
        gpu.lib().eco_array_urand_seed(it); gpu.lib().eco_array_grand_reset(); m.mutate()
        assert m.migrate() == o.migrate()
        assert m.condense() == o.condense()
        assert np.array_equal(m.density_array(), o.get_density_all()), it
        assert m.nsp() == o.nsp_per_cell().tolist()
        assert abs(m.connectivity() - o.connectivity()) <= 1e-15 * max(1.0, abs(o.connectivity()))


def test_checkpoint_restart_roundtrip(oracle_mod, gpu, tmp_path):
    """test/checkpoint-restart.sh:38-76: density, repro_rate, interaction.{diag,val,row,col} survive; and because
    Philox is stateless the restarted run continues bit-identically -- checked against the ORACLE run straight through
    (the reference excludes its RNG state, so its own restart is only statistically equivalent)."""
    o, m, grid = pair(oracle_mod, gpu, (8, 5), 37, 3)
    o.generate(3); o.migrate()
    m.generate(3); m.migrate()
    f = str(tmp_path / "ck.bin")
    m.checkpoint(f)
    m2 = gpu.SpatialModel(0)
    m2.restart(f)
    assert np.array_equal(m2.density_array(), o.get_density_all())
    assert m2.cell(3, 2).density() == m.cell(3, 2).density() == o.get_density(3 + 8 * 2).tolist()   # the geometry came back too
    assert m2.nsp() == m.nsp()
    for name in ("diag", "val", "row", "col"):
        assert getattr(m2.interaction, name)() == getattr(m.interaction, name)()
    assert m2.repro_rate() == m.repro_rate() and m2.tstep() == m.tstep() == int(o.get("tstep"))
    del m
    for _ in range(2):
        o.generate(4); m2.generate(4)
        assert m2.migrate() == o.migrate()
        assert np.array_equal(m2.density_array(), o.get_density_all())
    assert m2.condense() == o.condense()
    assert np.array_equal(m2.density_array(), o.get_density_all())
    # a panmictic checkpoint restores a panmictic model
    op, mp, _ = pair(oracle_mod, gpu, None, 20, 3)
    op.generate(5); mp.generate(5)
    mp.checkpoint(f)
    mp2 = gpu.PanmicticModel(0)
    mp2.restart(f)
    op.generate(5); mp2.generate(5)
    assert mp2.density() == op.get_density(0).tolist()


@pytest.mark.parametrize("grid,nsp,mut", [((300, 70), 60, 0.05), ((33, 5), 129, 0.2), ((70, 9), 1100, 0.01), ((9, 4), 2100, 0.02)])
def test_mutate_cells_records_in_reference_order(oracle_mod, gpu, grid, nsp, mut):
    """The cell side of mutate alone, with MANY events (>10^4 at the first shape -- far more than ModelData::mutate could
    digest): the device scan must deliver exactly the oracle's records, already ordered by (cell, slot) as the reference
    concatenates them (:233-240, gen_index), and subtract the mutants from their parents."""
    O = oracle_mod
    ncells = grid[0] * grid[1]
    p = synth(nsp, 2, 8, mut=mut)
    scal = dict(SCALARS, mut_max=mut)
    d = densities(ncells, nsp, 8)
    o = load_oracle(O, p, d, grid, scalars=scal); m = load_gpu(gpu, p, d, grid, scalars=scal)
    o.generate(2); m.generate(2)
    want = o.mutate_cells()
    got = m.mutate_cells()
    assert len(want) > (10 ** 4 if ncells > 20000 else 50)
    assert np.array_equal(got, want)
    assert np.array_equal(m.density_array(), o.get_density_all())


def test_errors_surface_as_exceptions(gpu):
    m = gpu.SpatialModel(0)
    m.species(range(4)); m.setGrid(3, 3)
    with pytest.raises(gpu.EcolabError):
        m.generate(1)                       # repro_rate / diag missing
    with pytest.raises(gpu.EcolabError):
        m.cell(5, 5).density([1, 2, 3, 4])  # outside the grid
    with pytest.raises(gpu.EcolabError):
        m.cell(0, 0).density([-1, 0, 0, 0])


def test_empty_model_is_harmless(gpu):
    m = gpu.SpatialModel(0)
    m.setGrid(4, 4)
    m.generate(1)
    assert m.migrate() == 0 and m.condense() == 0
    m.mutate()
    assert m.nsp() == [0] * 16


@pytest.mark.parametrize("nsp,conn,sigma", [(100, 3, 0.0), (64, 5, 1.5), (30, 0, 0.0), (200, 10, 3.0), (8, 7, 0.0)])
def test_random_interaction_matches_oracle(oracle_mod, gpu, nsp, conn, sigma):
    """ModelData::random_interaction (:43-47 -> sparse_mat::init_rand, include/sparse_mat.h:88-111): same libc stream,
    same pattern -- row sizes ~ N(conn, sigma), distinct columns, self-links dropped, val zero-filled."""
    O = oracle_mod
    p = synth(nsp, 0, 2)
    o = load_oracle(O, p, densities(1, nsp, 2), None)
    m = load_gpu(gpu, p, densities(1, nsp, 2), None)
    O.srand(77); O.reset_grand(); o.random_interaction(conn, sigma)
    gpu.lib().eco_array_urand_seed(77); gpu.lib().eco_array_grand_reset(); m.random_interaction(conn, sigma)
    for k in ("row", "col", "val"):
        assert list(getattr(m.interaction, k)()) == list(o.get("interaction." + k)), k
    row, col = np.array(m.interaction.row()), np.array(m.interaction.col())
    assert (row != col).all() and len(m.interaction.val) == len(row) and (np.array(m.interaction.val()) == 0).all()
    if len(row):
        assert row.max() < nsp and col.max() < nsp
        assert len(set(zip(row.tolist(), col.tolist()))) == len(row)      # distinct links


def test_panmictic_observables(oracle_mod, gpu):
    """PanmicticModel::lifetimes (:474-490), mutantConnectivity and extinctionConnectivity (:493-516) against the oracle
    (same entry-order sums: the values must be identical, not just close)."""
    O = oracle_mod
    nsp = 80
    p = synth(nsp, 4, 12, mut=2e-3)
    p["repro_rate"][::3] = -0.3                      # a third of the species dies out quickly
    d = densities(1, nsp, 12)
    scal = dict(SCALARS, mut_max=2e-3)
    o = load_oracle(O, p, d, None, scalars=scal); m = load_gpu(gpu, p, d, None, scalars=scal)
    seen = 0
    for it in range(8):
        before = len(m.species)
        o.generate(20); m.generate(20)
        O.srand(it); O.reset_grand(); o.mutate()
        gpu.lib().eco_array_urand_seed(it); gpu.lib().eco_array_grand_reset(); m.mutate()
        lt_g, lt_o = list(m.lifetimes()), o.lifetimes().tolist()
        assert lt_g == lt_o
        seen += len(lt_g)
        assert m.mutantConnectivity(before) == o.mutant_connectivity(before)
        assert m.extinctionConnectivity() == o.extinction_connectivity()
        if it == 0:
            assert o.extinction_connectivity() >= 0 and o.mutant_connectivity(before) >= 0
        assert m.condense() == o.condense()
        assert m.density() == o.get_density(0).tolist()
    assert seen > 0      # some species did live and die
