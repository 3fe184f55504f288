"""GPU: eco_species_totals_noalloc (include/ecolab_b200.h) against eco_species_totals, through the C-ABI.

The entry point has no counterpart in the reference (added for callers that may not call cudaMalloc): the totals kernel's
per-species accumulators sit in a static device array shared by every call on the GPU.  Integer work: the bar is bit-exact.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from helpers import densities, load_gpu, synth

pytestmark = pytest.mark.gpu

MAX_SPECIES = 16384   # ECO_NOALLOC_MAX_SPECIES


def _both(L, m, nsp):
    """(existing entry point, new entry point) on the model's current state, raw C-ABI calls."""
    a = np.full(nsp, -1, dtype=np.int64)
    b = np.full(nsp, -2, dtype=np.int64)
    assert L.eco_species_totals(m._h, a.ctypes.data_as(C.c_void_p), nsp) == 0
    assert L.eco_species_totals_noalloc(m._h, b.ctypes.data_as(C.c_void_p), nsp) == 0, L.eco_last_error()
    return a, b


def test_noalloc_matches_species_totals_on_consecutive_calls(gpu):
    L = gpu.lib()
    grid, nsp = (37, 11), 300
    m = load_gpu(gpu, synth(nsp, 3, 5, mig=0.07), densities(grid[0] * grid[1], nsp, 5), grid, seed=5)
    try:
        a1, b1 = _both(L, m, nsp)                       # call 1
        assert np.array_equal(a1, b1)
        assert np.array_equal(b1, m.density_array().astype(np.int64).sum(axis=0))
        m.generate(3); m.migrate()                      # new contents: call 2 must not see call 1's sums
        a2, b2 = _both(L, m, nsp)
        assert np.array_equal(a2, b2) and not np.array_equal(b1, b2)
        a3, b3 = _both(L, m, nsp)                       # unchanged state, straight after: same answer again
        assert np.array_equal(a3, b3) and np.array_equal(b3, b2)
    finally:
        m.close()


def test_noalloc_clears_what_the_last_call_left(gpu):
    """A large model, then a small one on the same GPU: the shared array still holds the large model's sums."""
    L = gpu.lib()
    big = load_gpu(gpu, synth(1500, 2, 7), densities(8 * 6, 1500, 7), (8, 6), seed=7)
    small = load_gpu(gpu, synth(40, 2, 8), densities(20 * 9, 40, 8), (20, 9), seed=8)
    pan = load_gpu(gpu, synth(64, 3, 9), densities(1, 64, 9), None, seed=9)
    try:
        for m, nsp in ((big, 1500), (small, 40), (pan, 64), (big, 1500), (small, 40)):
            a, b = _both(L, m, nsp)
            assert np.array_equal(a, b)
        assert np.array_equal(small.species_totals_noalloc(), small.density_array().astype(np.int64).sum(axis=0))
    finally:
        big.close(); small.close(); pan.close()


def test_noalloc_device_output_and_no_allocation(gpu):
    """`out` may be device memory, and the calls leave the device's free memory where it was."""
    import torch
    L = gpu.lib()
    grid, nsp = (64, 16), 500
    m = load_gpu(gpu, synth(nsp, 3, 11), densities(grid[0] * grid[1], nsp, 11), grid, seed=11)
    try:
        want, first = _both(L, m, nsp)                  # also the first launch of the kernel in this process
        assert np.array_equal(want, first)
        out_dev = torch.full((nsp,), -3, dtype=torch.int64, device="cuda:0")
        host = np.empty(nsp, dtype=np.int64)
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info(0)
        assert L.eco_species_totals_noalloc(m._h, C.c_void_p(out_dev.data_ptr()), nsp) == 0
        assert L.eco_species_totals_noalloc(m._h, host.ctypes.data_as(C.c_void_p), nsp) == 0
        free1, _ = torch.cuda.mem_get_info(0)
        assert free1 == free0
        assert np.array_equal(out_dev.cpu().numpy(), want) and np.array_equal(host, want)
    finally:
        m.close()


def test_noalloc_rejects_sizes_beyond_the_static_array(gpu):
    L = gpu.lib()
    m = load_gpu(gpu, synth(16, 2, 3), densities(4, 16, 3), (2, 2), seed=3)
    try:
        out = np.zeros(MAX_SPECIES + 1, dtype=np.int64)
        assert L.eco_species_totals_noalloc(m._h, out.ctypes.data_as(C.c_void_p), MAX_SPECIES + 1) != 0
        msg = L.eco_last_error().decode()
        assert "exceed" in msg and str(MAX_SPECIES) in msg
        assert L.eco_species_totals_noalloc(m._h, out.ctypes.data_as(C.c_void_p), 17) != 0      # not the model's count
        assert "differs" in L.eco_last_error().decode()
        a, b = _both(L, m, 16)                          # and the entry point still works afterwards
        assert np.array_equal(a, b)
    finally:
        m.close()
