"""ctypes binding of the C ABI declared in include/ecolab_b200.h.

The shared library (ecolab_b200/libecolab_b200.so) holds the sm_100a kernels and
the C++ host model.  There is no CPU fallback: if the library is missing or
cannot be loaded this module raises, and every call that fails raises
``EcolabError`` carrying ``eco_last_error()`` (the reference surfaces C++
exceptions as Python exceptions, include/error.h:24-49).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ECOLAB_B200_LIB") or os.path.join(_HERE, "libecolab_b200.so")   # override: profiling builds only


class EcolabError(RuntimeError):
    pass


class MutRecord(C.Structure):
    _fields_ = [("cell", C.c_uint32), ("slot", C.c_uint32), ("count", C.c_uint32)]


_vp, _cp, _sz = C.c_void_p, C.c_char_p, C.c_size_t
_u32, _u64, _i32, _dbl, _int = C.c_uint32, C.c_uint64, C.c_int32, C.c_double, C.c_int
_P = C.POINTER

# name -> argtypes; every entry point returns int status unless listed in _SPECIAL
SIGNATURES = {
    "eco_create": [_P(_vp), _int],
    "eco_destroy": [_vp],
    "eco_set_stream": [_vp, _vp],
    "eco_sync": [_vp],
    "eco_set_f64": [_vp, _cp, _vp, _sz],
    "eco_get_f64": [_vp, _cp, _vp, _sz, _P(_sz)],
    "eco_set_u32": [_vp, _cp, _vp, _sz],
    "eco_get_u32": [_vp, _cp, _vp, _sz, _P(_sz)],
    "eco_set_i32": [_vp, _cp, _vp, _sz],
    "eco_get_i32": [_vp, _cp, _vp, _sz, _P(_sz)],
    "eco_set_scalar": [_vp, _cp, _dbl],
    "eco_get_scalar": [_vp, _cp, _P(_dbl)],
    "eco_set_grid": [_vp, _sz, _sz],
    "eco_set_grid_slab": [_vp, _sz, _sz, _sz, _sz],
    "eco_set_density": [_vp, _sz, _vp, _sz],
    "eco_get_density": [_vp, _sz, _vp, _sz, _P(_sz)],
    "eco_set_density_all": [_vp, _vp, _sz, _sz],
    "eco_get_density_all": [_vp, _vp, _sz, _sz],
    "eco_get_density_rows": [_vp, _sz, _sz, _vp, _sz],
    "eco_set_density_rows": [_vp, _sz, _sz, _vp, _sz],
    "eco_set_density_all_device": [_vp, _vp, _sz, _sz],
    "eco_get_density_all_device": [_vp, _vp, _sz, _sz],
    "eco_fill_density": [_vp, _i32],
    "eco_fill_density_random": [_vp, _i32, _i32],
    "eco_get_salt": [_vp, _sz, _P(_dbl)],
    "eco_make_consistent": [_vp],
    "eco_seed": [_vp, _u32],
    "eco_generate": [_vp, C.c_uint],
    "eco_migrate": [_vp, _P(_u64)],
    "eco_step": [_vp, C.c_uint, _P(_u64)],
    "eco_step_async": [_vp, C.c_uint],
    "eco_migrate_wait": [_vp, _P(_u64)],
    "eco_step_host_begin": [_vp, _vp, _sz, _sz, C.c_uint],
    "eco_step_host": [_vp, _vp, _vp, _sz, _sz, C.c_uint, _P(_u64)],
    "eco_condense": [_vp, _P(C.c_uint)],
    "eco_mutate": [_vp],
    "eco_random_interaction": [_vp, C.c_uint, _dbl],
    "eco_connectivity": [_vp, _P(_dbl)],
    "eco_nsp_per_cell": [_vp, _vp, _sz],
    "eco_lifetimes": [_vp, _vp, _sz, _P(_sz)],
    "eco_mutant_connectivity": [_vp, _sz, _P(_dbl)],
    "eco_extinction_connectivity": [_vp, _P(_dbl)],
    "eco_species_totals": [_vp, _vp, _sz],
    "eco_species_totals_noalloc": [_vp, _vp, _sz],
    "eco_checkpoint": [_vp, _cp],
    "eco_restart": [_vp, _cp],
    "eco_migrate_local": [_vp, _P(_u64)],
    "eco_migrate_local_async": [_vp, _P(_vp)],
    "eco_alive_mask_device": [_vp, _P(_vp), _sz],
    "eco_halo_ptrs": [_vp, _P(_vp), _P(_vp), _P(_vp), _P(_vp), _P(_sz)],
    "eco_peer_export": [_vp, _vp, _sz],
    "eco_peer_connect": [_vp, _vp, _vp],
    "eco_peer_disconnect": [_vp],
    "eco_alive_mask": [_vp, _vp, _sz],
    "eco_condense_apply": [_vp, _vp, _sz, _P(C.c_uint)],
    "eco_mutate_cells": [_vp, _P(_sz)],
    "eco_mutate_records": [_vp, _P(MutRecord), _sz],
    "eco_mutate_model": [_vp, _vp, _sz],
    "eco_append_species": [_vp, _vp, _sz],
    "eco_set_tuning": [_vp, _cp, _int],
    "eco_pre_rounding": [_vp, _vp, _sz, _sz],
    "eco_kernel_times": [_vp, _vp, _sz, _P(_sz)],
}
_SPECIAL = {
    "eco_last_error": (_cp, []),
    "eco_version": (_cp, []),
    "eco_array_urand_seed": (None, [_int]),
    "eco_array_urand_rand": (None, [_vp, _sz]),
    "eco_array_grand_reset": (None, []),
}

_lib = None


def lib():
    """Load libecolab_b200.so (once).  Raises EcolabError if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise EcolabError(
                f"{LIB_PATH} not found: build it with `make -C ecolab_b200/csrc` "
                "(or __graft_entry__.build()); there is no CPU fallback")
        try:
            L = C.CDLL(LIB_PATH)
        except OSError as e:  # pragma: no cover
            raise EcolabError(f"cannot load {LIB_PATH}: {e}") from e
        for name, args in SIGNATURES.items():
            f = getattr(L, name)
            f.restype, f.argtypes = _int, args
        for name, (res, args) in _SPECIAL.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def check(status):
    if status != 0:
        raise EcolabError(lib().eco_last_error().decode("utf-8", "replace"))
