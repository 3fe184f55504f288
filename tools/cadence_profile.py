#!/usr/bin/env python
"""Where the scripts' cadence generate(100)+mutate()+migrate()+condense() spends its time (run under gpurun).
Splits mutate / condense into their C-ABI pieces and times each with a device synchronise on both sides."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import ecolab_b200 as E
from helpers import SCALARS, synth

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=64); ap.add_argument("--nsp", type=int, default=1000); ap.add_argument("--nnz", type=int, default=3)
ap.add_argument("--mutation", type=float, default=1e-5); ap.add_argument("--rounds", type=int, default=6)
ap.add_argument("--passes", type=int, default=1); ap.add_argument("--prealloc-gb", type=int, default=0, dest="prealloc")
a = ap.parse_args()
torch.cuda.set_device(0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
if a.prealloc:   # what bench.py leaves behind: large device buffers allocated and freed before this leg
    big = torch.empty(a.prealloc << 30, dtype=torch.uint8, device="cuda"); big.fill_(1); torch.cuda.synchronize(); del big
    torch.cuda.empty_cache()
for pass_ in range(a.passes):
    print("pass", pass_, flush=True)
    m = E.SpatialModel(0, None); m.set_stream(st)
    p = synth(a.nsp, a.nnz, 10, mig=0.1)
    for k, v in SCALARS.items(): getattr(m, k)(v)
    for k, v in p.items():
        (getattr(m.interaction, k.split(".")[1]) if k.startswith("interaction.") else getattr(m, k))(v)
    m.setGrid(a.n, a.n); m.seed(10)
    m.mut_max(a.mutation); m.mutation(np.full(a.nsp, a.mutation))
    m.fill_density(100)
    E.lib().eco_array_urand_seed(10)
    def T(f):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r = f(); torch.cuda.synchronize(); return (time.perf_counter() - t0) * 1e3, r
    rows = []
    for r in range(a.rounds):
        d = {"nsp_before": len(m.species)}
        d["generate100"], _ = T(lambda: m.generate(100))
        d["mutate_cells"], rec = T(m.mutate_cells)
        parents = np.repeat(rec[:, 1], rec[:, 2]).astype(np.uint32); cells = np.repeat(rec[:, 0], rec[:, 2]).astype(np.uint32)
        d["mutants"] = int(len(parents))
        d["mutate_model"], _ = T(lambda: m.mutate_model(parents))
        d["append_species"], _ = T(lambda: m.append_species(cells) if len(cells) else None)
        d["migrate"], _ = T(m.migrate)
        d["alive_mask"], mask = T(m.alive_mask)
        d["condense_apply"], rem = T(lambda: m.condense_apply(mask))
        d["removed"] = int(rem)
        d["generate1_after(upload)"], _ = T(lambda: m.generate(1))
        d["nnz"] = len(m.interaction.val)
        rows.append({k: (round(v, 3) if isinstance(v, float) else v) for k, v in d.items()})
        print(rows[-1], flush=True)
