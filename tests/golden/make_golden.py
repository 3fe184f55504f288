"""Generate the committed golden fixtures (tests/golden/*.npz) from the CPU oracle in Philox mode.

    python tests/golden/make_golden.py        (from the repo root)

The reference itself cannot run here (missing classdesc submodule), so the fixtures come from the
restatement, which test_oracle_golden.py pins to the reference's own golden vector first.  Each case
is a short scripted run; the fixture stores the densities (and scalars) after every stage.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from helpers import densities, load_oracle, synth   # noqa: E402

# name -> (grid, nsp, nnz/row, seed, script)
CASES = {
    "panmictic_100sp": (None, 100, 3, 1, ["g1", "g5", "c", "g100"]),
    "grid8x5_37sp": ((8, 5), 37, 3, 2, ["g1", "m", "g3", "m", "c", "g1", "m"]),
    "grid33x3_64sp_k16": ((33, 3), 64, 16, 3, ["g2", "m", "g1", "m"]),
    "grid64x2_9sp_dense": ((64, 2), 9, 8, 4, ["g1", "m", "c", "g4"]),
}


def run_script(model_ops, script):
    out = {}
    for i, op in enumerate(script):
        if op[0] == "g":
            model_ops["generate"](int(op[1:]))
        elif op == "m":
            out[f"{i}_moved"] = np.array([model_ops["migrate"]()], dtype=np.int64)
        elif op == "c":
            out[f"{i}_removed"] = np.array([model_ops["condense"]()], dtype=np.int64)
        out[f"{i}_{op}"] = model_ops["density"]()
    return out


def case_inputs(name):
    grid, nsp, k, seed, script = CASES[name]
    ncells = grid[0] * grid[1] if grid else 1
    p = synth(nsp, k, seed)
    d = densities(ncells, nsp, seed)
    d[:, nsp // 3] = 0          # one species extinct everywhere -> condense has work
    return grid, p, d, script


def run_case(O, name):
    grid, p, d, script = case_inputs(name)
    o = load_oracle(O, p, d, grid, seed=10 + len(name))
    ops = {"generate": o.generate, "migrate": o.migrate, "condense": o.condense,
           "density": (lambda: o.get_density_all()) if grid else (lambda: np.array([o.get_density(0)], dtype=np.int32))}
    return run_script(ops, script)


if __name__ == "__main__":
    from oracle import oracle as O
    O.build()
    for name in CASES:
        out = run_case(O, name)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, {k: v.shape for k, v in out.items()})
