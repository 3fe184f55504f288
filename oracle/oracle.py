"""ctypes wrapper around oracle/libecolab_oracle.so -- TEST INFRASTRUCTURE ONLY.

The oracle is the CPU restatement of the reference's hot path (see the header of
oracle/ecolab_oracle.cc).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product package
(ecolab_b200/) must never do so.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libecolab_oracle.so")

RNG_MT19937 = 0
RNG_PHILOX = 1
ST_GENERATE, ST_MUTATE, ST_SALT = 0, 1, 2


def build(force=False):
    src = os.path.join(_HERE, "ecolab_oracle.cc")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        vp, cp, sz, u32, u64, dbl = C.c_void_p, C.c_char_p, C.c_size_t, C.c_uint32, C.c_uint64, C.c_double
        sig = {
            "orc_create": (vp, []),
            "orc_destroy": (None, [vp]),
            "orc_set_rng": (None, [vp, C.c_int, u32]),
            "orc_seed": (None, [vp, u32]),
            "orc_reset_grand": (None, []),
            "orc_set_threads": (None, [C.c_int]),
            "orc_max_threads": (C.c_int, []),
            "orc_set_scalar": (C.c_int, [vp, cp, dbl]),
            "orc_get_scalar": (dbl, [vp, cp]),
            "orc_set_f64": (C.c_int, [vp, cp, vp, sz]),
            "orc_get_f64": (C.c_long, [vp, cp, vp, sz]),
            "orc_set_u32": (C.c_int, [vp, cp, vp, sz]),
            "orc_get_u32": (C.c_long, [vp, cp, vp, sz]),
            "orc_set_species": (None, [vp, vp, sz]),
            "orc_get_species": (C.c_long, [vp, vp, sz]),
            "orc_set_grid": (None, [vp, sz, sz]),
            "orc_set_density": (None, [vp, sz, vp, sz]),
            "orc_fill_density_random": (None, [vp, C.c_int32, C.c_int32]),
            "orc_set_grid_slab": (None, [vp, sz, sz, sz, sz]),
            "orc_migrate_local": (u64, [vp]),
            "orc_set_ghost": (None, [vp, C.c_int, vp, sz]),
            "orc_alive_mask": (None, [vp, vp]),
            "orc_condense_apply": (C.c_uint, [vp, vp]),
            "orc_mutate_cells": (sz, [vp]),
            "orc_mutate_records": (None, [vp, vp]),
            "orc_append_species": (None, [vp, vp, sz]),
            "orc_get_density": (C.c_long, [vp, sz, vp, sz]),
            "orc_set_density_all": (None, [vp, vp, sz]),
            "orc_get_density_all": (None, [vp, vp, sz]),
            "orc_get_salt": (dbl, [vp, sz]),
            "orc_make_consistent": (None, [vp]),
            "orc_generate": (None, [vp, C.c_uint]),
            "orc_migrate": (u64, [vp]),
            "orc_condense": (C.c_uint, [vp]),
            "orc_mutate": (None, [vp]),
            "orc_mutate_model": (None, [vp, vp, sz]),
            "orc_random_interaction": (None, [vp, C.c_uint, dbl]),
            "orc_connectivity": (dbl, [vp]),
            "orc_mutant_connectivity": (dbl, [vp, sz]),
            "orc_extinction_connectivity": (dbl, [vp]),
            "orc_lifetimes": (C.c_long, [vp, vp, sz]),
            "orc_nsp_per_cell": (None, [vp, vp]),
            "orc_get_pre": (C.c_long, [vp, vp, sz]),
            "orc_philox": (u32, [u32, u32, u32, u32, u32, u64]),
            "orc_philox_raw": (None, [vp, vp, vp]),
            "orc_round": (C.c_int, [dbl, u32]),
        }
        for name, (res, args) in sig.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


F64 = ("create", "repro_rate", "mutation", "migration", "interaction.diag", "interaction.val")
U32 = ("interaction.row", "interaction.col")
SCALARS = ("sp_sep", "mut_max", "repro_min", "repro_max", "odiag_min", "odiag_max", "gen_bias",
           "tstep", "last_mut_tstep", "last_mig_tstep", "fixMutation", "fixMigration")


class Oracle:
    """One model instance (panmictic until set_grid is called)."""

    def __init__(self, rng=RNG_PHILOX, seed=0):
        self.L = lib()
        self.h = self.L.orc_create()
        self.L.orc_set_rng(self.h, rng, seed)

    def __del__(self):
        try:
            self.L.orc_destroy(self.h)
        except Exception:
            pass

    # --- members -----------------------------------------------------------
    def set(self, name, value):
        if name in F64:
            a = np.ascontiguousarray(value, dtype=np.float64)
            assert self.L.orc_set_f64(self.h, name.encode(), _ptr(a), a.size) == 0
        elif name in U32:
            a = np.ascontiguousarray(value, dtype=np.uint32)
            assert self.L.orc_set_u32(self.h, name.encode(), _ptr(a), a.size) == 0
        elif name == "species":
            a = np.ascontiguousarray(value, dtype=np.int32)
            self.L.orc_set_species(self.h, _ptr(a), a.size)
        else:
            assert self.L.orc_set_scalar(self.h, name.encode(), float(value)) == 0, name

    def get(self, name):
        if name in F64:
            n = self.L.orc_get_f64(self.h, name.encode(), None, 0)
            a = np.empty(n, dtype=np.float64)
            self.L.orc_get_f64(self.h, name.encode(), _ptr(a), n)
            return a
        if name in U32:
            n = self.L.orc_get_u32(self.h, name.encode(), None, 0)
            a = np.empty(n, dtype=np.uint32)
            self.L.orc_get_u32(self.h, name.encode(), _ptr(a), n)
            return a
        if name == "species":
            n = self.L.orc_get_species(self.h, None, 0)
            a = np.empty(n, dtype=np.int32)
            self.L.orc_get_species(self.h, _ptr(a), n)
            return a
        return self.L.orc_get_scalar(self.h, name.encode())

    @property
    def nsp(self):
        return int(self.get("nsp"))

    @property
    def ncells(self):
        return int(self.get("ncells"))

    # --- grid / density ----------------------------------------------------
    def set_grid(self, nx, ny):
        self.L.orc_set_grid(self.h, nx, ny)

    # --- row-slab halves (multi-rank CPU tests) -------------------------------
    def set_grid_slab(self, nx, ny, y0, nyl):
        self.L.orc_set_grid_slab(self.h, nx, ny, y0, nyl)

    def migrate_local(self):
        return int(self.L.orc_migrate_local(self.h))

    def set_ghost(self, bottom, rows):
        a = np.ascontiguousarray(rows, dtype=np.int32)
        self.L.orc_set_ghost(self.h, int(bottom), _ptr(a), a.size)

    def alive_mask(self):
        a = np.zeros(self.nsp, dtype=np.uint8)
        self.L.orc_alive_mask(self.h, _ptr(a))
        return a

    def condense_apply(self, mask):
        a = np.ascontiguousarray(mask, dtype=np.uint8)
        return int(self.L.orc_condense_apply(self.h, _ptr(a)))

    def mutate_cells(self):
        n = self.L.orc_mutate_cells(self.h)
        a = np.zeros((n, 3), dtype=np.uint32)
        if n:
            self.L.orc_mutate_records(self.h, _ptr(a))
        return a

    def append_species(self, cells):
        a = np.ascontiguousarray(cells, dtype=np.uint32)
        self.L.orc_append_species(self.h, _ptr(a), a.size)

    def fill_density_random(self, lo, hi):
        self.L.orc_fill_density_random(self.h, lo, hi)

    def set_density(self, cell, d):
        a = np.ascontiguousarray(d, dtype=np.int32)
        self.L.orc_set_density(self.h, cell, _ptr(a), a.size)

    def get_density(self, cell):
        n = self.L.orc_get_density(self.h, cell, None, 0)
        a = np.empty(n, dtype=np.int32)
        self.L.orc_get_density(self.h, cell, _ptr(a), n)
        return a

    def set_density_all(self, d):
        a = np.ascontiguousarray(d, dtype=np.int32)
        assert a.ndim == 2 and a.shape[0] == self.ncells
        self.L.orc_set_density_all(self.h, _ptr(a), a.shape[1])

    def get_density_all(self):
        n = self.L.orc_get_density(self.h, 0, None, 0)
        a = np.empty((self.ncells, n), dtype=np.int32)
        self.L.orc_get_density_all(self.h, _ptr(a), n)
        return a

    def salt(self, cell):
        return self.L.orc_get_salt(self.h, cell)

    # --- the hot path --------------------------------------------------------
    def seed(self, x):
        self.L.orc_seed(self.h, x)

    def make_consistent(self):
        self.L.orc_make_consistent(self.h)

    def generate(self, niter=1):
        self.L.orc_generate(self.h, niter)

    def migrate(self):
        return int(self.L.orc_migrate(self.h))

    def condense(self):
        return int(self.L.orc_condense(self.h))

    def mutate(self):
        self.L.orc_mutate(self.h)

    def mutate_model(self, parents):
        a = np.ascontiguousarray(parents, dtype=np.uint32)
        self.L.orc_mutate_model(self.h, _ptr(a), a.size)

    def random_interaction(self, conn, sigma):
        self.L.orc_random_interaction(self.h, conn, sigma)

    def connectivity(self):
        return self.L.orc_connectivity(self.h)

    def mutant_connectivity(self, before_nsp):
        return self.L.orc_mutant_connectivity(self.h, int(before_nsp))

    def extinction_connectivity(self):
        return self.L.orc_extinction_connectivity(self.h)

    def lifetimes(self):
        a = np.empty(max(self.nsp, 1), dtype=np.float64)
        n = self.L.orc_lifetimes(self.h, _ptr(a), a.size)
        return a[:n]

    def nsp_per_cell(self):
        a = np.empty(self.ncells, dtype=np.uint32)
        self.L.orc_nsp_per_cell(self.h, _ptr(a))
        return a

    def pre_rounding(self):
        n = self.L.orc_get_pre(self.h, None, 0)
        a = np.empty(n, dtype=np.float64)
        self.L.orc_get_pre(self.h, _ptr(a), n)
        return a


def philox(seed, stage, x, y, slot, t):
    """The shared stream's draw for cell (x, y), species slot `slot`, timestep t."""
    return int(lib().orc_philox(seed, stage, x, y, slot, t))


def philox_raw(ctr, key):
    """Plain Philox4x32-10 block (known-answer tests against the Random123 vectors)."""
    c = np.array(ctr, dtype=np.uint32); k = np.array(key, dtype=np.uint32); o = np.zeros(4, dtype=np.uint32)
    lib().orc_philox_raw(_ptr(c), _ptr(k), _ptr(o))
    return [int(v) for v in o]


def round_with_draw(x, u):
    return int(lib().orc_round(float(x), int(u)))


def reset_grand():
    lib().orc_reset_grand()


def set_threads(n):
    lib().orc_set_threads(int(n))


def max_threads():
    return int(lib().orc_max_threads())


_libc = C.CDLL(None)


def srand(s):
    """libc srand (the reference's array_urand.seed, include/random_basic.h:15)."""
    _libc.srand(C.c_uint(s))
