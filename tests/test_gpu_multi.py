"""GPU, world_size 2 (and 4/8 when present): the partitioned model over NCCL against the whole-grid oracle."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("halo", ["peer", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_partitioned_model_matches_oracle(world, halo):
    """Row-slab partition over `world` GPUs against the single-process oracle of the whole grid: generate, migrate (halo
    rows read out of the neighbours' memory over NVLink, or exchanged through NCCL), condense, mutate incl. the new species'
    parameters, buffer re-allocation, non-blocking steps, the scripts' SPMD cell access."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, ECOLAB_B200_HALO=halo)
    for attempt in range(3):   # the free port found by _port() can be taken again before torchrun binds it
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
               "--master-port", str(_port()), os.path.join(HERE, "multi_gpu_worker.py")]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        if "EADDRINUSE" not in r.stderr:
            break
    assert r.returncode == 0 and f"MULTI_GPU_OK {world} {halo}" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("halo", ["peer", "nccl"])
def test_partitioned_model_retiles_on_every_rank(halo):
    """560 species of which 61 are extinct: condense() crosses the 512-row threshold, every rank re-lays its slab out on
    32-cell tiles (same decision everywhere), the halo reconnects to the new buffers, and the run stays the oracle's."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, ECOLAB_B200_HALO=halo, ECO_MULTI_NSP="560", ECO_MULTI_DEAD="60")
    for attempt in range(3):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
               "--master-port", str(_port()), os.path.join(HERE, "multi_gpu_worker.py")]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        if "EADDRINUSE" not in r.stderr:
            break
    assert r.returncode == 0 and f"MULTI_GPU_OK 2 {halo}" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
