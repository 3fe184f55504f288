"""CPU: host-side logic of the model mirror and of the multi-GPU driver (gloo, world_size 2)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_slab_bounds_cover_the_grid():
    from ecolab_b200 import slab_bounds
    for ny in (1, 2, 7, 64, 4096):
        for w in (1, 2, 3, 4, 8):
            if ny < w:
                continue
            b = [slab_bounds(ny, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == ny
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [y1 - y0 for y0, y1 in b]
            assert min(sizes) >= 1 and max(sizes) - min(sizes) <= 1


def test_record_merge_and_expand():
    from ecolab_b200.spatial import expand_records, merge_records
    a = np.array([[3, 1, 2], [3, 5, 1]], dtype=np.uint32)
    b = np.array([[9, 0, 1]], dtype=np.uint32)
    m = merge_records([b, np.zeros((0, 3), np.uint32), a])
    assert m.tolist() == [[3, 1, 2], [3, 5, 1], [9, 0, 1]]
    parents, cells = expand_records(m)
    assert parents.tolist() == [1, 1, 5, 0] and cells.tolist() == [3, 3, 3, 9]
    assert merge_records([]).shape == (0, 3)


def test_ecolist_protocol():
    from ecolab_b200 import EcoList
    v = EcoList([1, 1, 1])
    assert v == [1, 1, 1] and v() == [1, 1, 1] and v[0] == 1 and [x * 2 for x in v] == [2, 2, 2]


def test_compat_modules_import_headless():
    import ecolab_b200.compat  # noqa: F401
    import ecolab
    import GUI
    import plot
    import ecolab_model
    assert ecolab.myid() == 0 and ecolab.nprocs() == 1 and ecolab.workGroupSize(128) == 128
    GUI.statusBar.configure(text="t=0")
    plot.plot("No. species", 0, 3)
    GUI.gui(lambda: None, None)
    assert hasattr(ecolab_model, "spatial_ecolab") and hasattr(ecolab_model, "panmictic_ecolab")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch.distributed as dist
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from ecolab_b200.spatial import Comm
        from slab_emulation import run_distributed_against_oracle
        comm = Comm()
        # 1. the small collectives
        mask = np.array([1, 0, 0, 1] if rank == 0 else [0, 0, 1, 1], dtype=np.uint8)
        assert comm.allreduce_max_u8(mask).tolist() == [1, 0, 1, 1]
        assert comm.allreduce_sum(10 + rank) == sum(10 + r for r in range(world))
        rec = np.array([[rank * 100 + 1, 2, 1]] * (rank + 1), dtype=np.uint32)
        got = comm.allgather_records(rec)
        assert [len(g) for g in got] == [r + 1 for r in range(world)] and got[1][0].tolist() == [101, 2, 1]
        assert [len(g) for g in comm.allgather_records(np.zeros((0, 3), np.uint32))] == [0] * world
        t = comm.bcast_arrays({"a": np.arange(3)} if rank == 0 else None)
        assert t["a"].tolist() == [0, 1, 2]
        # 2. the whole distributed driver (slab partition, halo, mask reduce, mutant gather) against the oracle
        run_distributed_against_oracle(comm)
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:   # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))


@pytest.mark.timeout(180)
def test_distributed_driver_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=170) for _ in ps]
    for p in ps:
        p.join(10)
    for r, msg in res:
        assert msg == "ok", f"rank {r}: {msg}"


def test_array_urand_is_libc_rand():
    """array_urand (src/random.cc:57: rand()/RAND_MAX) is restated inside the library so that the model owns the state: it
    must be libc's own rand() stream, however the draws are grouped."""
    import ctypes
    import numpy as np
    from ecolab_b200 import _cabi
    if not os.path.exists(_cabi.LIB_PATH):
        pytest.skip("library not built")
    L = ctypes.CDLL(_cabi.LIB_PATH)          # no device needed: the generator is host state
    L.eco_array_urand_seed.argtypes = [ctypes.c_int]; L.eco_array_urand_seed.restype = None
    L.eco_array_urand_rand.argtypes = [ctypes.c_void_p, ctypes.c_size_t]; L.eco_array_urand_rand.restype = None
    libc = ctypes.CDLL("libc.so.6")
    libc.rand.restype = ctypes.c_int

    def draw(n):
        a = np.empty(n, dtype=np.float64)
        L.eco_array_urand_rand(a.ctypes.data_as(ctypes.c_void_p), n)
        return a

    for seed in (1, 10, 12345):
        libc.srand(seed)
        want = np.array([libc.rand() for _ in range(5000)], dtype=np.float64) / 2147483647.0
        L.eco_array_urand_seed(seed)
        one = np.concatenate([draw(1) for _ in range(5000)])
        assert np.array_equal(one, want)
        L.eco_array_urand_seed(seed)
        assert np.array_equal(draw(5000), want)                      # one block
        L.eco_array_urand_seed(seed)
        sizes, got = [1, 8, 3, 191, 192, 193, 2, 577, 31, 64, 7, 1000, 9, 5], []
        for n in sizes:                                              # singles and blocks interleaved: the state carries over
            got.append(draw(n))
        got = np.concatenate(got)
        assert np.array_equal(got, want[:len(got)])
