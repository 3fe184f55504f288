"""Shadow model for running the reference's driver scripts UNCHANGED against the oracle (tests only).

`install(kind)` puts a module named `ecolab_model` into sys.modules whose `spatial_ecolab` / `panmictic_ecolab` objects
forward every call to the product's compat objects (ecolab_b200.compat -> C ABI -> CUDA) and mirror it into a CPU
oracle instance fed the same inputs.  After every hot-path call (generate / migrate / condense / mutate) the return value
and every density are compared with the oracle's.  A budget of generate() calls bounds scripts that loop for millions of
timesteps: the next generate() raises StopScript (a BaseException, so no script-level `except` swallows it).

What is mirrored rather than recomputed: the draws of ModelData::mutate and random_interaction come from the process-wide
libc rand() stream, which product and oracle would otherwise consume in turn; the shadow lets the product draw, checks the
deterministic cell side of mutate (which cells speciate, how many mutants, the parents' loss) against the oracle, and
copies the new species' parameters into the oracle.  Their distribution is tested elsewhere (test_gpu_parity.py).
"""
import sys
import types

import numpy as np

F64 = ("create", "repro_rate", "mutation", "migration")
INTER_F64 = ("diag", "val")
INTER_U32 = ("row", "col")
SCALARS = ("sp_sep", "mut_max", "repro_min", "repro_max", "odiag_min", "odiag_max", "gen_bias", "tstep", "last_mut_tstep",
           "last_mig_tstep", "fixMutation", "fixMigration")


class StopScript(BaseException):
    pass


class _Member:
    def __init__(self, sh, gmember, oname):
        self._sh, self._g, self._o = sh, gmember, oname

    def __call__(self, *a):
        if a:
            v = self._g(*a)
            self._sh._mirror(self._o)
            return v
        return self._g()

    def __len__(self):
        return len(self._g)

    def __getitem__(self, i):
        return self._g[i]

    def __iter__(self):
        return iter(self._g)

    @property
    def _properties(self):
        return self._g._properties


class _Interaction:
    def __init__(self, sh):
        g = sh._g.interaction
        for n in INTER_F64 + INTER_U32:
            setattr(self, n, _Member(sh, getattr(g, n), "interaction." + n))
        self.rowsz, self.colsz = g.rowsz, g.colsz


class _Cell:
    def __init__(self, sh, i, j):
        self._sh = sh
        g = sh._g.cell(i, j)
        cid = int(i) + sh._nx * int(j)

        class _Density(_Member):
            def __call__(s, *a):   # noqa: N805
                if a:
                    v = g.density(*a)
                    sh._o.set_density(cid, np.asarray(v, dtype=np.int32))
                    return v
                v = g.density()
                assert list(v) == sh._o.get_density(cid).tolist(), f"cell {cid} differs from the oracle"
                return v

        self.density = _Density(sh, g.density, None)
        self.salt, self.id = g.salt, g.id

    def nsp(self):
        return sum(1 for v in self.density() if v != 0)


class Shadow:
    def __init__(self, gpu_model, O, spatial, budget=None):
        self._g, self._O, self._spatial, self._budget = gpu_model, O, spatial, budget
        self._o = O.Oracle(rng=O.RNG_PHILOX, seed=0)
        self._nx = self._ny = 1
        self.generates = self.checks = 0
        for n in SCALARS:
            self._mirror(n)
        for n in F64:
            setattr(self, n, _Member(self, getattr(gpu_model, n), n))
        self.species = _Member(self, gpu_model.species, "species")
        self.interaction = _Interaction(self)
        for n in SCALARS:
            setattr(self, n, _Member(self, getattr(gpu_model, n), n))
        if not spatial:
            sh = self

            class _Density(_Member):
                def __call__(s, *a):   # noqa: N805
                    if a:
                        v = gpu_model.density(*a)
                        sh._o.set_density(0, np.asarray(v, dtype=np.int32))
                        return v
                    return gpu_model.density()

            self.density = _Density(self, gpu_model.density, None)

    # ---- mirroring --------------------------------------------------------------------------
    def _gval(self, name):
        obj = self._g
        for part in name.split("."):
            obj = getattr(obj, part)
        return obj()

    def _mirror(self, name):
        v = self._gval(name)
        self._o.set(name, float(v) if name in SCALARS else np.asarray(v))

    def _gdens(self):
        if self._spatial:
            return self._g.density_array()
        return np.array([self._g.density()], dtype=np.int32)

    def _odens(self):
        if self._spatial:
            return self._o.get_density_all()
        return np.array([self._o.get_density(0)], dtype=np.int32)

    def _pad_oracle(self):
        """The product pads every cell to the model's species count with zero densities (what the comment at
        SpatialModel::makeConsistent, models/ecolab_model.cc:727, asks for); models/migration_test.py populates ONE cell
        and leaves the others empty, so the oracle's cells are padded the same way before any hot-path call."""
        if not self._spatial:
            return
        nc = self._nx * self._ny
        lens = [len(self._o.get_density(c)) for c in range(nc)]
        ns = max([self._o.nsp] + lens)
        for c in range(nc):
            if lens[c] < ns:
                d = self._o.get_density(c)
                self._o.set_density(c, np.concatenate([d, np.zeros(ns - len(d), dtype=np.int32)]))

    def _check(self, what):
        self.checks += 1
        g, o = self._gdens(), self._odens()
        assert g.shape == o.shape and np.array_equal(g, o), f"{what}: densities differ from the oracle (tstep {self._g.tstep()})"
        assert self._g.tstep() == int(self._o.get("tstep"))

    # ---- geometry ----------------------------------------------------------------------------
    def setGrid(self, nx, ny):
        self._g.setGrid(nx, ny)
        self._o.set_grid(int(nx), int(ny))
        self._nx, self._ny = int(nx), int(ny)

    def cell(self, i, j):
        return _Cell(self, i, j)

    # ---- the hot path -------------------------------------------------------------------------
    def seed(self, x):
        self._g.seed(x)
        self._o.seed(int(x) & 0xFFFFFFFF)

    def makeConsistent(self):
        self._g.makeConsistent()
        self._o.make_consistent()
        for n in ("species",) + F64:
            self._mirror(n)

    def generate(self, niter=1):
        if self._budget is not None and self.generates >= self._budget:
            raise StopScript()
        self.generates += 1
        self._pad_oracle()
        self._g.generate(niter)
        self._o.generate(int(niter))
        self._check("generate")

    def migrate(self):
        self._pad_oracle()
        r, ro = self._g.migrate(), self._o.migrate()
        assert r == ro, f"migrate returned {r}, the oracle {ro}"
        self._check("migrate")
        return r

    def condense(self):
        self._pad_oracle()
        r, ro = self._g.condense(), self._o.condense()
        assert r == ro, f"condense removed {r}, the oracle {ro}"
        self._check("condense")
        for n in ("species", "repro_rate", "interaction.val", "interaction.row", "interaction.col"):
            assert np.array_equal(np.asarray(self._gval(n)), self._o.get(n)), n
        return r

    def mutate(self):
        self._pad_oracle()
        n0 = len(self._g.species)
        rec = self._o.mutate_cells()                      # deterministic: the stage-1 Philox stream
        born = int(rec[:, 2].sum()) if len(rec) else 0
        self._g.mutate()
        assert len(self._g.species) == n0 + born, "number of mutants differs from the oracle's speciation records"
        if born:
            for n in ("species",) + F64 + tuple("interaction." + k for k in INTER_F64 + INTER_U32):
                self._mirror(n)
            cells = np.repeat(rec[:, 0], rec[:, 2].astype(np.int64)).astype(np.uint32)
            self._o.append_species(cells)
        self._mirror("last_mut_tstep")
        self._check("mutate")

    def random_interaction(self, conn, sigma):
        self._g.random_interaction(conn, sigma)
        for n in ("interaction.row", "interaction.col", "interaction.val"):
            self._mirror(n)

    # ---- observables ----------------------------------------------------------------------------
    def nsp(self):
        v = self._g.nsp()
        if self._spatial:
            self._pad_oracle()
            assert list(v) == self._o.nsp_per_cell().tolist()
        return v

    def connectivity(self):
        v = self._g.connectivity()
        assert v == self._o.connectivity()
        return v

    def lifetimes(self):
        v = self._g.lifetimes()
        assert list(v) == self._o.lifetimes().tolist()
        for n in ("create",):
            self._mirror(n)
        return v

    def __getattr__(self, name):   # partitionObjects, gather, syncThreads, checkpoint, ...: nothing to mirror
        return getattr(self._g, name)


def install(O, budget=None):
    """Fresh product objects behind shadows under the module name the scripts import.  Returns the module."""
    import ecolab_b200.compat  # noqa: F401  (puts the shim directory on sys.path)
    for name in ("ecolab_model",):
        sys.modules.pop(name, None)
    import ecolab_model as real
    mod = types.ModuleType("ecolab_model")
    mod.spatial_ecolab = Shadow(real._make_spatial(), O, True, budget)
    mod.panmictic_ecolab = Shadow(real._make_panmictic(), O, False, budget)
    sys.modules["ecolab_model"] = mod
    return mod


def uninstall():
    sys.modules.pop("ecolab_model", None)
