"""Script-level acceptance, dynamic half: the call sequences of the reference's driver scripts
(models/migration_test.py, spatial_ecolab.py / steady_migration.py, SAR.py, panmictic_ecolab.py, pred-prey.py), restated
here and driven through the SAME module surface a reference user imports -- `from ecolab_model import spatial_ecolab as
ecolab`, `from ecolab import array_urand, myid`, `from plot import plot`, `from GUI import statusBar` -- on the GPU, with
the CPU oracle fed the same inputs as the checker.  The reference tree itself does not exist on the GPU box; that every
member the real scripts touch exists is checked statically in tests/test_script_surface.py.
"""
import random

import numpy as np
import pytest

from helpers import SCALARS, load_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture()
def mods(gpu):
    import ecolab_b200.compat  # noqa: F401
    import ecolab
    import ecolab_model
    import GUI
    import plot
    return ecolab_model, ecolab, GUI, plot


def _rand_list(n, lo, hi):
    return [random.random() * (hi - lo) + lo for _ in range(n)]


def _params_of(model):
    """What the script fed the model, read back through the getters (the oracle gets exactly these)."""
    p = {k: np.array(getattr(model, k)()) for k in ("species", "repro_rate", "mutation", "migration", "create")}
    for k in ("diag", "val", "row", "col"):
        p["interaction." + k] = np.array(getattr(model.interaction, k)())
    p["species"] = p["species"].astype(np.int32)
    p["interaction.row"] = p["interaction.row"].astype(np.uint32)
    p["interaction.col"] = p["interaction.col"].astype(np.uint32)
    return p


def _scalars_of(model):
    return {k: float(getattr(model, k)()) for k in SCALARS}


def _dens(model, nx, ny):
    return np.array([model.cell(i, j).density() for j in range(ny) for i in range(nx)], dtype=np.int32)


def test_migration_test_py(oracle_mod, mods):
    """models/migration_test.py: 10x10 grid, 100 species, one populated cell, migration 0.5 (capped at 1/maxNbrs)."""
    em, eco, GUI, plot = mods
    m = em.spatial_ecolab
    eco.array_urand.seed(10)
    nsp, nx, ny = 100, 10, 10
    m.species(range(nsp))
    m.setGrid(nx, ny)
    m.partitionObjects()
    m.cell(4, 4).density(nsp * [100])
    m.migration(nsp * [0.5])
    m.makeConsistent()
    m.tstep(0); m.last_mig_tstep(0)
    m.seed(10)
    o = load_oracle(oracle_mod, _params_of(m), _dens(m, nx, ny), (nx, ny), scalars=_scalars_of(m))
    for step in range(6):
        moved = m.migrate()
        m.tstep(m.tstep() + 1)
        assert moved == o.migrate()
        o.set("tstep", o.get("tstep") + 1)
        d = _dens(m, nx, ny)
        assert np.array_equal(d, o.get_density_all())
        assert (d.sum(axis=0) == 100).all() and (d >= 0).all()          # per-species conservation
        per_cell = m.nsp()()
        assert len(per_cell) == nx * ny and per_cell == [int((r != 0).sum()) for r in d]
        plot.plot("No. species by cell", m.tstep(), per_cell)
        plot.plot("Density(4,4)", m.tstep(), m.cell(4, 4).density(), pens=m.species())
    assert m.cell(4, 2).density()[0] > 0 and m.cell(3, 3).density()[0] > 0   # the front has reached two cells away


def _init_spatial(m, eco, nsp, nx, ny, mut_max, migration, odiag=1e-5, seed=10):
    """The initialisation block shared by spatial_ecolab.py:17-56, steady_migration.py:17-52 and SAR.py:47-76."""
    random.seed(seed)
    eco.array_urand.seed(seed + eco.myid())
    m.tstep(0); m.last_mig_tstep(0); m.last_mut_tstep(0)
    m.repro_min(-0.1); m.repro_max(0.1)
    m.odiag_min(-odiag); m.odiag_max(odiag)
    m.mut_max(mut_max); m.sp_sep(0.1); m.gen_bias(0.0)
    m.species(range(nsp))
    m.create(nsp * [0])
    m.setGrid(nx, ny)
    m.partitionObjects()
    for i in range(nx):
        for j in range(ny):
            m.cell(i, j).density(nsp * [100])
    m.repro_rate(_rand_list(nsp, m.repro_min(), m.repro_max()))
    m.interaction.diag(_rand_list(nsp, -1e-3, -1e-3))
    m.random_interaction(3, 0)
    m.interaction.val(_rand_list(len(m.interaction.val), m.odiag_min(), m.odiag_max()))
    m.mutation(nsp * [mut_max])
    m.migration(nsp * [migration])
    m.makeConsistent()
    m.syncThreads()
    m.seed(10)           # the checker's Philox key (pred-prey.py calls seed(10) itself; the spatial scripts keep the default)


def test_steady_migration_py(oracle_mod, mods, gpu):
    """steady_migration.py / spatial_ecolab.py: 2x2 grid, 100 species, generate(100); mutate(); migrate(); condense()."""
    em, eco, GUI, plot = mods
    m = em.spatial_ecolab
    O = oracle_mod
    nsp, nx, ny = 100, 2, 2
    m.fixMigration(True); m.fixMutation(False)
    _init_spatial(m, eco, nsp, nx, ny, mut_max=1e-3, migration=1e-3, odiag=1e-4)
    assert m.nsp()() == [nsp] * (nx * ny)
    o = load_oracle(O, _params_of(m), _dens(m, nx, ny), (nx, ny), scalars=_scalars_of(m))
    o.set("fixMigration", 1)
    saved = m.migration()
    assert saved == nsp * [1e-3] and [x * 1000 for x in saved][0] == pytest.approx(1.0)
    extinctions = migrations = 0
    for it in range(5):
        m.generate(100); o.generate(100)
        gpu.lib().eco_array_urand_seed(it); gpu.lib().eco_array_grand_reset(); m.mutate()
        O.srand(it); O.reset_grand(); o.mutate()
        mv = m.migrate()
        assert mv == o.migrate()
        ex = m.condense()
        assert ex == o.condense()
        migrations += mv; extinctions += ex
        d = _dens(m, nx, ny)
        assert np.array_equal(d, o.get_density_all()) and (d >= 0).all()
        n_now = len(m.species)
        assert n_now == d.shape[1] == len(m.repro_rate()) == len(m.interaction.diag())
        assert abs(m.connectivity() - o.connectivity()) <= 1e-15 * max(1.0, abs(o.connectivity()))
        GUI.statusBar.configure(text=f"t={m.tstep()} nsp:{n_now}")
        plot.plot("No. species by cell", m.tstep(), m.nsp()())
    assert m.tstep() == 500 and extinctions >= 0 and migrations >= 0


@pytest.mark.parametrize("area,nsp,mut_max,migration", [(1, 30, 1e-5, 0.0), (4, 100, 1e-3, 1e-3), (6, 30, 1e-4, 1e-4), (2, 200, 1e-3, 1e-5)])
def test_sar_py_reuses_one_model_object(oracle_mod, mods, gpu, area, nsp, mut_max, migration):
    """SAR.py:39-82: the SAME global model object is re-initialised for every (area, rate, species count) combination --
    new grid, new species count, tstep back to 0 -- and stepped generate(100); mutate(); migrate(); condense(); gather()."""
    em, eco, GUI, plot = mods
    m = em.spatial_ecolab
    O = oracle_mod
    m.fixMutation(True); m.fixMigration(True)
    nx = int(np.sqrt(area)); ny = area // nx
    _init_spatial(m, eco, nsp, nx, ny, mut_max=mut_max, migration=migration, seed=area * 7 + nsp)
    o = load_oracle(O, _params_of(m), _dens(m, nx, ny), (nx, ny), scalars=_scalars_of(m))
    o.set("fixMigration", 1); o.set("fixMutation", 1)
    for it in range(4):
        m.generate(100); o.generate(100)
        gpu.lib().eco_array_urand_seed(100 + it); gpu.lib().eco_array_grand_reset(); m.mutate()
        O.srand(100 + it); O.reset_grand(); o.mutate()
        assert m.migrate() == o.migrate()
        assert m.condense() == o.condense()
        m.gather()
        assert np.array_equal(_dens(m, nx, ny), o.get_density_all())
    av = lambda x: sum(x) / len(x) if len(x) > 0 else 0   # noqa: E731
    assert av(m.mutation()) == pytest.approx(mut_max) and av(m.migration()) == pytest.approx(migration)   # fixed rates
    assert len(m.species) == o.get_density_all().shape[1]


def test_panmictic_ecolab_py(oracle_mod, mods, gpu):
    """panmictic_ecolab.py: one cell, 100 species, generate(100); mutate(); condense(); connectivity each step."""
    em, eco, GUI, plot = mods
    m = em.panmictic_ecolab
    O = oracle_mod
    random.seed(3)
    eco.array_urand.seed(10)
    nsp = 100
    m.tstep(0); m.last_mut_tstep(0)
    m.repro_min(-0.1); m.repro_max(0.1); m.odiag_min(-1e-5); m.odiag_max(1e-5); m.mut_max(1e-4); m.sp_sep(0.1); m.gen_bias(0.0)
    m.species(range(nsp))
    m.density(nsp * [100])
    m.create(nsp * [0])
    m.repro_rate(_rand_list(nsp, m.repro_min(), m.repro_max()))
    m.interaction.diag(_rand_list(nsp, -1e-3, -1e-3))
    m.random_interaction(3, 0)
    m.interaction.val(_rand_list(len(m.interaction.val), m.odiag_min(), m.odiag_max()))
    m.mutation(nsp * [m.mut_max()])
    m.migration(nsp * [0])
    m.seed(10)
    p = _params_of(m)
    o = load_oracle(O, p, np.array([m.density()], dtype=np.int32), None, scalars=_scalars_of(m))
    for it in range(6):
        m.generate(100); o.generate(100)
        gpu.lib().eco_array_urand_seed(it); gpu.lib().eco_array_grand_reset(); m.mutate()
        O.srand(it); O.reset_grand(); o.mutate()
        assert m.condense() == o.condense()
        assert m.density() == o.get_density(0).tolist()
        assert len(m.species) == len(m.density())
        plot.plot("No. species", m.tstep(), len(m.species))
        assert abs(m.connectivity() - o.connectivity()) <= 1e-15 * max(1.0, abs(o.connectivity()))
    assert m.tstep() == 600


def test_pred_prey_py(oracle_mod, mods):
    """pred-prey.py: two species with ids [1, 2], everything set member by member, generate() with its default niter,
    density()._properties read every step; no migration array, no mutate, no condense."""
    em, eco, GUI, plot = mods
    m = em.panmictic_ecolab
    m.tstep(0)
    m.species([1, 2])
    m.density([100, 100])
    m.create([0, 0])
    m.repro_rate([.1, -.1])
    m.interaction.diag([-.0001, -1e-5])
    m.interaction.val([-0.001, 0.001])
    m.interaction.row([0, 1])
    m.interaction.col([1, 0])
    m.mutation([.01, .01])
    m.sp_sep(.1); m.repro_min(-.1); m.repro_max(.1); m.odiag_min(-1e-3); m.odiag_max(1e-3); m.mut_max(.01)
    m.seed(10)
    p = _params_of(m)
    p["migration"] = np.zeros(2)
    o = load_oracle(oracle_mod, p, np.array([m.density()], dtype=np.int32), None, scalars=_scalars_of(m))
    for _ in range(60):
        m.generate(); o.generate(1)
        density = m.density()._properties
        GUI.statusBar.configure(text=f"t={m.tstep()} n={density}")
        plot.plot("density", m.tstep(), density)
        assert m.density() == o.get_density(0).tolist()
    assert m.tstep() == 60 and m.species() == [1, 2]


def test_fragmentation_pump_py(oracle_mod, mods, gpu):
    """fragmentation_pump.py:63-80: the migration rates are switched off (all 0) and back on (1e-3) between epochs while the
    generate(100); mutate(); migrate(); condense() loop keeps running."""
    em, eco, GUI, plot = mods
    m = em.spatial_ecolab
    O = oracle_mod
    nsp, nx, ny = 60, 3, 2
    m.fixMigration(True); m.fixMutation(False)
    _init_spatial(m, eco, nsp, nx, ny, mut_max=1e-3, migration=1e-3, odiag=1e-4, seed=4)
    o = load_oracle(O, _params_of(m), _dens(m, nx, ny), (nx, ny), scalars=_scalars_of(m))
    o.set("fixMigration", 1)
    saved = None
    for it in range(6):
        m.generate(100); o.generate(100)
        gpu.lib().eco_array_urand_seed(it); gpu.lib().eco_array_grand_reset(); m.mutate()
        O.srand(it); O.reset_grand(); o.mutate()
        if it == 2:                                    # half epoch: fragment the landscape
            saved = m.migration()
            m.migration(len(m.species) * [0]); o.set("migration", np.zeros(len(m.species)))
        if it == 4:                                    # epoch: reconnect it
            m.migration(len(m.species) * [1e-3]); o.set("migration", np.full(len(m.species), 1e-3))
        mv = m.migrate()
        assert mv == o.migrate()
        if it in (2, 3):
            assert mv == 0
        assert m.condense() == o.condense()
        assert np.array_equal(_dens(m, nx, ny), o.get_density_all())
    assert saved is not None and len(saved) > 0 and all(x == 1e-3 for x in saved)


def test_spatial_ecolab_sh(mods):
    """test/00/spatial_ecolab.sh:18-77 as a fresh process would run it: parameters BEFORE setGrid, generate() with its
    default argument, every cell changes, migrate() conserves every species' total."""
    em, eco, GUI, plot = mods
    m = em._Lazy(em._make_spatial)          # a fresh process-lifetime object, as in a new interpreter
    random.seed(1)
    nsp = 30
    m.species(range(nsp))
    m.repro_min(-.1); m.repro_max(.1); m.odiag_min(-1e-5); m.odiag_max(1e-5)
    m.repro_rate(_rand_list(nsp, m.repro_min(), m.repro_max()))
    m.interaction.diag(_rand_list(nsp, -1e-3, -1e-3))
    m.random_interaction(3, 0)
    m.interaction.val(_rand_list(len(m.interaction.val), m.odiag_min(), m.odiag_max()))
    m.mut_max(1e-4); m.sp_sep(0.1)
    m.mutation(nsp * [m.mut_max()])
    m.migration(nsp * [1e-5])
    m.setGrid(2, 2)
    initd = nsp * [100]
    for i, j in ((0, 0), (0, 1), (1, 0), (1, 1)):
        m.cell(i, j).density(initd)
    m.distributeObjects()
    m.generate()
    m.gather()
    for i, j in ((0, 0), (0, 1), (1, 0), (1, 1)):
        assert m.cell(i, j).density() != initd
    n0 = nsp * [0]
    for i in range(2):
        for j in range(2):
            d = m.cell(i, j).density()
            for k in range(nsp):
                n0[k] += d[k]
    m.tstep(m.tstep() + 999)                # so that dt*migration really moves individuals (1e-5 * 1000 * ~100 >= 1)
    moved = m.migrate()
    n1 = nsp * [0]
    for i in range(2):
        for j in range(2):
            d = m.cell(i, j).density()
            for k in range(nsp):
                n1[k] += d[k]
    assert n0 == n1 and moved >= 0


def test_checkpoint_restart_sh(gpu, tmp_path):
    """test/checkpoint-restart.sh:38-76 (a TCL script in the reference; the same statements through the Python surface):
    set the panmictic model, checkpoint, wipe two members, restart, every member is back."""
    ecolab = gpu.PanmicticModel(0)      # a fresh model, as the freshly started `models/ecolab` interpreter of the script has
    density, repro_rate, diag = [100, 100], [.1, -.1], [-.0001, 0]
    val, row, col = [-0.001, 0.001], [0, 1], [1, 0]
    ecolab.density(density)
    ecolab.repro_rate(repro_rate)
    ecolab.interaction.diag(diag)
    ecolab.interaction.val(val)
    ecolab.interaction.row(row)
    ecolab.interaction.col(col)
    f = str(tmp_path / "test.ckpt")
    ecolab.checkpoint(f)
    ecolab.repro_rate([])
    ecolab.interaction.val([])
    ecolab.restart(f)
    assert ecolab.density() == density and ecolab.repro_rate() == repro_rate
    assert ecolab.interaction.diag() == diag and ecolab.interaction.val() == val
    assert ecolab.interaction.row() == row and ecolab.interaction.col() == col
