"""`ecolab`: the small part of the reference's base module the model scripts use
(src/ecolab.cc:225-266): array_urand / array_grand, myid, nprocs, device,
workGroupSize, ecolabHome, and inert Parallel / Plot / register hooks."""
import os as _os


class _Urand:
    """array_urand (src/arrays.cc:24): libc rand() behind the model's parameter mutation."""

    def seed(self, s):
        from ecolab_b200 import lib
        lib().eco_array_urand_seed(int(s))

    def rand(self):
        """urand::rand (src/random.cc:57): the next rand()/RAND_MAX of the stream the model's parameter mutation draws from."""
        import ctypes
        from ecolab_b200 import lib
        v = ctypes.c_double()
        lib().eco_array_urand_rand(ctypes.byref(v), 1)
        return v.value


class _Grand:
    """array_grand (src/arrays.cc:25): the running-CLT gaussian generator."""

    def reset(self):
        from ecolab_b200 import lib
        lib().eco_array_grand_reset()


array_urand = _Urand()
array_grand = _Grand()
_wgs = [256]


def myid():
    return int(_os.environ.get("RANK", "0"))


def nprocs():
    return int(_os.environ.get("WORLD_SIZE", "1"))


def device():
    from ecolab_b200 import lib
    return lib().eco_version().decode()


def workGroupSize(n=None):
    """src/ecolab.cc:232,238.  Kept for the scripts; launch shapes are chosen by the engine."""
    if n is not None:
        _wgs[0] = int(n)
    return _wgs[0]


def ecolabHome():
    return _os.path.dirname(_os.path.abspath(__file__))


def registerParallel():
    return None


class Parallel:
    """src/ecolab.cc:130-229 re-broadcasts method names to MPI workers; here every rank runs the script (SPMD)."""

    def __init__(self, target=None):
        self._t = target

    def __call__(self, method, *args):
        return getattr(self._t, method)(*args)

    def exit(self):
        return None


class Plot:
    def __getattr__(self, n):
        return lambda *a, **k: None


def register(obj, name):
    globals()[name] = obj
