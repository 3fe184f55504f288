"""Script-level acceptance for real: the reference's OWN driver scripts -- verbatim copies under
tests/ref_scripts/ (see its README) -- executed unchanged with runpy through `ecolab_b200.compat` on the GPU.

Single GPU: every script runs against a shadow (tests/shadow.py) that mirrors each call into the CPU oracle and compares
all densities and return values after every generate / migrate / condense / mutate; scripts that loop for millions of
timesteps are stopped after a budget of generate() calls.  Two GPUs (torchrun): the spatial scripts run unchanged on a
row-slab partition and must leave exactly the state of the single-GPU run.
"""
import json
import os
import runpy
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
SCRIPTS = os.path.join(HERE, "ref_scripts")

# (script, generate() budget or None = run to the end, kind)
CASES = [
    ("test00_predator_prey.py", None),        # 1000 x generate() + 10 x condense, writes out.dat
    ("test00_condense.py", None),             # its own assert: density == [1, 1, 1]
    ("test00_spatial_ecolab.py", None),       # its own asserts: every cell changes; migrate conserves every species
    ("migration_test.py", None),              # gui(step) -> ECOLAB_STEPS steps of migrate from a point source
    ("pred-prey.py", None),
    ("panmictic_ecolab.py", None),            # gui(step): ECOLAB_STEPS x (generate(100) + mutate + condense)
    ("spatial_ecolab.py", 25),                # stops inside its timeit(stepImpl, number=100)
    ("steady_migration.py", 25),              # generate(100)+mutate+migrate+condense, writes steady.dat
    ("fragmentation_pump.py", 25),            # the same with the migration pump, writes pump.dat
    ("SAR.py", 12),                           # first ensemble member (1 x 1 grid, 30 species)
    ("SAR1.py", 12),                          # the variant that takes the migration rate from its command line
]
ARGS = {"SAR1.py": ["1e-3"]}


@pytest.mark.parametrize("script,budget", CASES)
def test_reference_script_unchanged_against_oracle(oracle_mod, gpu, tmp_path, monkeypatch, script, budget):
    import shadow
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("ECOLAB_STEPS", "6")
    monkeypatch.setattr(sys, "argv", [script] + ARGS.get(script, []))
    mod = shadow.install(oracle_mod, budget)
    try:
        try:
            runpy.run_path(os.path.join(SCRIPTS, script), run_name="__main__")
            stopped = False
        except shadow.StopScript:
            stopped = True
    finally:
        shadow.uninstall()
    sh = mod.panmictic_ecolab if mod.panmictic_ecolab.checks else mod.spatial_ecolab
    assert stopped == (budget is not None)
    if script == "test00_predator_prey.py":
        rows = [list(map(int, l.split())) for l in open("out.dat")]
        assert len(rows) == 10 and all(len(r) == 2 and min(r) >= 0 for r in rows)
        assert sh.generates == 1000 and rows[-1] == list(sh._g.density())
    elif script == "test00_condense.py":
        assert list(sh._g.density()) == [1, 1, 1]
    elif script in ("steady_migration.py", "fragmentation_pump.py"):
        head = open("steady.dat" if script.startswith("steady") else "pump.dat").readline()
        assert head.startswith("#tstep")
    if script not in ("test00_condense.py",):
        assert sh.checks >= (budget or 1)            # the oracle really followed the run
        assert sh._g.tstep() == int(sh._o.get("tstep"))


def _port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("script,budget", [("test00_spatial_ecolab.py", 99), ("spatial_ecolab.py", 6), ("steady_migration.py", 8),
                                           ("fragmentation_pump.py", 8), ("migration_test.py", 0)])
def test_reference_script_unchanged_on_two_ranks(tmp_path, script, budget):
    """`torchrun --nproc-per-node 2 script.py` semantics: every rank executes the same script; assignments to cells of
    another rank are dropped, nsp()() returns all cells, cell reads come from the owner, gather() synchronises -- and the
    final state is bit-identical to the single-GPU run of the same script."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    worker = os.path.join(HERE, "ref_script_worker.py")
    outs = []
    for world in (1, 2):
        out = str(tmp_path / f"fp{world}.json")
        cwd = tmp_path / f"w{world}"
        cwd.mkdir()
        if world == 1:
            cmd = [sys.executable, worker, os.path.join(SCRIPTS, script), str(budget), out]
        else:
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                   "--master-port", str(_port()), worker, os.path.join(SCRIPTS, script), str(budget), out]
        for attempt in range(3):   # the free port found by _port() can be taken again before torchrun binds it
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=str(cwd))
            if "EADDRINUSE" not in r.stderr:
                break
            if world > 1:
                cmd[cmd.index("--master-port") + 1] = str(_port())
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
        outs.append(json.load(open(out)))
    a, b = outs
    # everything but the species LABELS: a new species is labelled sp_cntr + k*nprocs (models/ecolab_model.cc:463-464), so the
    # labels of a 2-rank run advance twice as fast -- by the reference's own rule; they must still be unique
    for k in ("generates", "tstep", "nsp", "density", "individuals", "nsp_per_cell", "repro_rate", "val"):
        assert a[k] == b[k], (k, a[k], b[k])
    assert a["species_unique"] and b["species_unique"]
    assert b["world"] == 2
