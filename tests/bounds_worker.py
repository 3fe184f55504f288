"""Runs a spread of operations through WHATEVER library ECOLAB_B200_LIB points at and reports the bounds-assert counters
(tests/test_gpu_bounds.py loads libecolab_b200_bounds.so, the -DECO_BOUNDS build of the same sources)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ecolab_b200 as E   # noqa: E402
from helpers import SCALARS, densities, load_gpu, synth   # noqa: E402

out = {"cases": 0, "bad": 0, "sites": [], "selftest": None}
rng = np.random.default_rng(1)
shapes = [((1, 1), 7, 2), ((33, 3), 64, 16), ((70, 9), 129, 3), ((250, 3), 1000, 3), ((17, 4), 1030, 5), ((40, 7), 1700, 5),
          ((7, 41), 2100, 1), ((6, 20), 3700, 1), ((130, 70), 300, 9), ((1030, 20), 60, 2), ((31, 1), 5, 0), ((2500, 1), 1000, 3)]
for k, (grid, nsp, nnz) in enumerate(shapes):
    mut = min(0.02, 40.0 / (grid[0] * grid[1] * nsp))     # a few dozen mutants per mutate(): capacity growth, not an explosion
    p = synth(nsp, nnz, 50 + k, mig=0.11, mut=mut)
    scal = dict(SCALARS, mut_max=mut)
    d = densities(grid[0] * grid[1], nsp, 50 + k)
    d[:, nsp // 2] = 0
    m = load_gpu(E, p, d, grid, seed=k, scalars=scal)
    if k == 0:
        out["selftest"] = int(m._scalar("bounds_selftest"))
    if k % 3 == 1:
        m.set_tuning("mig_rows", int(rng.choice([1, 3, 16])))
    if k % 4 == 2:
        m.set_tuning("gen_nobsmem", 1)
    m.generate(1); m.migrate(); m.generate(3); m.mutate(); m.migrate(); m.condense(); m.generate(1)
    buf = m.density_array()
    m.step_host(buf, buf, 1)
    m.fill_density_random(0, 50); m.step(2); m.nsp(); m.species_totals()
    bad = int(m._scalar("bounds_bad"))
    out["cases"] += 1
    out["bad"] += bad
    if bad:
        out["sites"].append([k, int(m._scalar("bounds_site"))])
    m.close()
print("BOUNDS_RESULT " + json.dumps(out))
