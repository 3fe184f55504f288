"""Host-side mirror of the reference's scripted model objects.

The reference exports its model as two process-lifetime, classdesc-reflected
globals, ``panmictic_ecolab`` and ``spatial_ecolab`` (models/ecolab_model.cc:32-41),
through an attribute-call protocol pinned by its driver scripts
(models/spatial_ecolab.py, models/panmictic_ecolab.py, test/00/*.sh):

    m.member()            -> value (a list for array members)
    m.member(v)           -> set (accepts any iterable / range) and return the new value
    m.interaction.diag(v) -> nested objects by attribute
    len(m.species)        -> element count
    m.generate(100); m.condense() -> int; m.migrate() -> int; m.cell(i, j).density(v); m.nsp()()

This module reproduces that surface over the C ABI (include/ecolab_b200.h).  All
model state lives in the C++ object behind the handle (host parameter arrays,
device-resident densities); getters copy out, setters copy in, exactly the
ownership rule of the reference wrapper.  Nothing here computes: without the CUDA
library every call raises ``EcolabError``.
"""
import ctypes as C

import numpy as np

from . import _cabi
from ._cabi import EcolabError, check

_F64 = ("create", "repro_rate", "mutation", "migration")
_F64_I = ("diag", "val")
_U32_I = ("row", "col")
_SCALARS_F = ("sp_sep", "mut_max", "repro_min", "repro_max", "odiag_min", "odiag_max", "gen_bias")
_SCALARS_I = ("tstep", "last_mut_tstep", "last_mig_tstep")
_SCALARS_B = ("fixMutation", "fixMigration")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class EcoList(list):
    """What an array getter returns: a plain list (so ``== [1, 1, 1]`` holds,
    test/00/condense.sh:20) that is also callable, because the reference hands
    back wrapped arrays for method results (``ecolab.nsp()()``,
    models/spatial_ecolab.py:46)."""

    @property
    def _properties(self):
        """The plain value behind the wrapper (`ecolab.density()._properties`, models/pred-prey.py:23)."""
        return list(self)

    def __call__(self):
        return list(self)


class _Member:
    """One reflected data member: callable getter/setter (CppWrapperType in the reference)."""

    def __init__(self, get, set_=None, length=None, name=""):
        self._get, self._set, self._length, self._name = get, set_, length, name

    def __call__(self, *args):
        if args:
            if self._set is None:
                raise EcolabError(f"{self._name} is read-only")
            self._set(args[0] if len(args) == 1 else list(args))
        return self._get()

    def __len__(self):
        if self._length is not None:
            return self._length()
        return len(self._get())

    def __getitem__(self, i):
        return self._get()[i]

    def __iter__(self):
        return iter(self._get())

    @property
    def _properties(self):
        return self._get()

    def _list(self):
        return []

    def __repr__(self):
        return f"<ecolab member {self._name}>"


class _Interaction:
    """ModelData::interaction (sparse_mat, include/sparse_mat.h:20-32)."""

    def __init__(self, model):
        m = model
        for n in _F64_I:
            setattr(self, n, m._array_member("interaction." + n, "f64"))
        for n in _U32_I:
            setattr(self, n, m._array_member("interaction." + n, "u32"))
        self.rowsz = _Member(lambda: int(m._scalar("nsp")), name="interaction.rowsz")
        self.colsz = _Member(lambda: int(m._scalar("nsp")), name="interaction.colsz")

    def _list(self):
        return ["diag", "val", "row", "col", "rowsz", "colsz"]


class _Cell:
    """EcolabCell (models/ecolab_model.h:123-126): density, salt, id of one grid cell."""

    def __init__(self, model, cid):
        self._m, self._id = model, cid
        self.density = _Member(lambda: model._get_density(cid), lambda v: model._set_density(cid, v),
                               name=f"cell[{cid}].density")
        self.salt = _Member(lambda: model._salt(cid), name="salt")
        self.id = _Member(lambda: cid, name="id")

    def nsp(self):
        return sum(1 for v in self.density() if v != 0)

    def _list(self):
        return ["density", "salt", "id"]


class EcolabModelBase:
    """State and methods shared by PanmicticModel and SpatialModel (ModelData, models/ecolab_model.h:39-68)."""

    def __init__(self, device=0):
        self._L = _cabi.lib()
        h = C.c_void_p()
        check(self._L.eco_create(C.byref(h), int(device)))
        self._h = h
        self._device = int(device)
        for n in _F64:
            setattr(self, n, self._array_member(n, "f64"))
        self.species = self._array_member("species", "i32")
        self.interaction = _Interaction(self)
        for n in _SCALARS_F:
            setattr(self, n, self._scalar_member(n, float))
        for n in _SCALARS_I:
            setattr(self, n, self._scalar_member(n, int))
        for n in _SCALARS_B:
            setattr(self, n, self._scalar_member(n, bool))

    def close(self):
        """Destroy the C++ model now (device buffers included).  The members are closures over `self`, so the object sits in
        a reference cycle and `del m` alone frees the device memory only at the next garbage collection."""
        if getattr(self, "_h", None):
            self._L.eco_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing ----------------------------------------------------------
    def _scalar(self, name):
        out = C.c_double()
        check(self._L.eco_get_scalar(self._h, name.encode(), C.byref(out)))
        return out.value

    def _set_scalar(self, name, v):
        check(self._L.eco_set_scalar(self._h, name.encode(), float(v)))

    def _scalar_member(self, name, typ):
        return _Member(lambda: typ(self._scalar(name)), lambda v: self._set_scalar(name, v), name=name)

    def _get_array(self, name, kind):
        fn = getattr(self._L, "eco_get_" + kind)
        n = C.c_size_t()
        check(fn(self._h, name.encode(), None, 0, C.byref(n)))
        dt = {"f64": np.float64, "u32": np.uint32, "i32": np.int32}[kind]
        a = np.empty(n.value, dtype=dt)
        if n.value:
            check(fn(self._h, name.encode(), _ptr(a), a.size, C.byref(n)))
        return a

    def _set_array(self, name, kind, v):
        dt = {"f64": np.float64, "u32": np.uint32, "i32": np.int32}[kind]
        a = np.ascontiguousarray(list(v) if not isinstance(v, np.ndarray) else v, dtype=dt)
        check(getattr(self._L, "eco_set_" + kind)(self._h, name.encode(), _ptr(a), a.size))

    def _array_len(self, name, kind):
        n = C.c_size_t()
        check(getattr(self._L, "eco_get_" + kind)(self._h, name.encode(), None, 0, C.byref(n)))
        return n.value

    def _array_member(self, name, kind):
        return _Member(lambda: EcoList(self._get_array(name, kind).tolist()),
                       lambda v: self._set_array(name, kind, v),
                       lambda: self._array_len(name, kind), name=name)

    def _get_density(self, cell):
        n = C.c_size_t()
        check(self._L.eco_get_density(self._h, cell, None, 0, C.byref(n)))
        a = np.empty(n.value, dtype=np.int32)
        if n.value:
            check(self._L.eco_get_density(self._h, cell, _ptr(a), a.size, C.byref(n)))
        return EcoList(a.tolist())

    def _set_density(self, cell, v):
        a = np.ascontiguousarray(list(v) if not isinstance(v, np.ndarray) else v, dtype=np.int32)
        check(self._L.eco_set_density(self._h, cell, _ptr(a), a.size))

    def _salt(self, cell):
        out = C.c_double()
        check(self._L.eco_get_salt(self._h, cell, C.byref(out)))
        return out.value

    # ---- numpy-speed accessors (not part of the reference surface) -----------
    def density_array(self):
        """All owned cells as an int32 array [cell][nsp] (ascending cell id)."""
        nc, ns = int(self._scalar("ncells_local")), int(self._scalar("nsp"))
        a = np.empty((nc, ns), dtype=np.int32)
        if a.size:
            check(self._L.eco_get_density_all(self._h, _ptr(a), nc, ns))
        return a

    def set_density_array(self, a):
        a = np.ascontiguousarray(a, dtype=np.int32)
        check(self._L.eco_set_density_all(self._h, _ptr(a), a.shape[0], a.shape[1]))

    def species_totals(self):
        ns = int(self._scalar("nsp"))
        a = np.empty(ns, dtype=np.int64)
        if ns:
            check(self._L.eco_species_totals(self._h, _ptr(a), ns))
        return a

    def species_totals_noalloc(self, out=None):
        """species_totals() through eco_species_totals_noalloc: static device scratch, no cudaMalloc on the call
        (at most ECO_NOALLOC_MAX_SPECIES = 16384 species; no counterpart in the reference).  `out`: a caller-owned
        int64 numpy array of nsp entries, allocated here when omitted."""
        ns = int(self._scalar("nsp"))
        a = np.empty(ns, dtype=np.int64) if out is None else out
        if ns:
            check(self._L.eco_species_totals_noalloc(self._h, _ptr(a), ns))
        return a

    def pre_rounding(self):
        nc, ns = int(self._scalar("ncells_local")), int(self._scalar("nsp"))
        a = np.empty((nc, ns), dtype=np.float64)
        check(self._L.eco_pre_rounding(self._h, _ptr(a), nc, ns))
        return a

    def fill_density(self, v):
        check(self._L.eco_fill_density(self._h, int(v)))

    def fill_density_random(self, lo, hi):
        """n = lo + philox(seed, stage 3, cell, slot, t=0) % (hi-lo+1) on every owned cell (synthetic initial state)."""
        check(self._L.eco_fill_density_random(self._h, int(lo), int(hi)))

    def launches(self):
        return int(self._scalar("launches"))

    def set_tuning(self, knob, value):
        check(self._L.eco_set_tuning(self._h, knob.encode(), int(value)))

    def kernel_times(self):
        a = np.zeros(16, dtype=np.float64)
        n = C.c_size_t()
        check(self._L.eco_kernel_times(self._h, _ptr(a), a.size, C.byref(n)))
        names = ("generate", "migrate", "step", "dense_pack", "dense_mma")
        return {names[i]: (a[2 * i], int(a[2 * i + 1])) for i in range(n.value // 2)}

    # ---- the reflected methods ---------------------------------------------------
    def makeConsistent(self):
        check(self._L.eco_make_consistent(self._h))

    def seed(self, x):
        check(self._L.eco_seed(self._h, int(x) & 0xFFFFFFFF))

    def generate(self, niter=1):
        check(self._L.eco_generate(self._h, int(niter)))

    def condense(self):
        r = C.c_uint()
        check(self._L.eco_condense(self._h, C.byref(r)))
        return r.value

    def mutate(self):
        check(self._L.eco_mutate(self._h))

    def random_interaction(self, conn, sigma):
        check(self._L.eco_random_interaction(self._h, int(conn), float(sigma)))

    def connectivity(self):
        out = C.c_double()
        check(self._L.eco_connectivity(self._h, C.byref(out)))
        return out.value

    def complexity(self):
        raise EcolabError("complexity() (nauty network complexity, src/netcomplexity.cc) is outside the "
                          "population-update path and is not provided")

    def checkpoint(self, file):
        check(self._L.eco_checkpoint(self._h, str(file).encode()))

    def restart(self, file):
        check(self._L.eco_restart(self._h, str(file).encode()))

    def syncThreads(self):
        check(self._L.eco_sync(self._h))

    def _list(self):
        return sorted(n for n in dir(self) if not n.startswith("_"))

    @property
    def _properties(self):
        return {n: getattr(self, n)() for n in _SCALARS_F + _SCALARS_I + _SCALARS_B}


class PanmicticModel(EcolabModelBase):
    """panmictic_ecolab (models/ecolab_model.h:105-121): the model object is its single cell."""

    def __init__(self, device=0):
        super().__init__(device)
        self.density = _Member(lambda: self._get_density(0), lambda v: self._set_density(0, v), name="density")
        self.salt = _Member(lambda: self._salt(0), name="salt")

    def lifetimes(self):
        ns = max(int(self._scalar("nsp")), 1)
        a = np.empty(ns, dtype=np.float64)
        n = C.c_size_t()
        check(self._L.eco_lifetimes(self._h, _ptr(a), a.size, C.byref(n)))
        return EcoList(a[: n.value].tolist())

    def mutantConnectivity(self, before_nsp):
        out = C.c_double()
        check(self._L.eco_mutant_connectivity(self._h, int(before_nsp), C.byref(out)))
        return out.value

    def extinctionConnectivity(self):
        out = C.c_double()
        check(self._L.eco_extinction_connectivity(self._h, C.byref(out)))
        return out.value

    def updateConnectionPlot(self):   # Cairo drawing (models/ecolab_model.cc:519-578): no graphics stack, no-op
        return None

    def nsp(self):
        return sum(1 for v in self.density() if v != 0)
