#!/usr/bin/env python
"""bench_aux.py -- the kernels either side of generate/migrate on C3 (1024 x 1024 cells x 1024 species):
condense (alive mask + compaction), mutate (speciation scan + append), observables (nsp per cell, species totals),
generate(100) (the scripts' cadence, on-chip multi-step kernel).  Wall time of the synchronous API call
(perf_counter around the call, device idle before) and the achieved fraction of the measured HBM peak for the
algorithmic bytes of each pass.  Run it plain, and under
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv
for the per-kernel device times and DRAM traffic (profiles/r1_aux_kernels.txt).

  python tools/bench_aux.py [--n 1024] [--nsp 1024] > gpurun_out/aux.json
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1024)
    ap.add_argument("--nsp", type=int, default=1024)
    ap.add_argument("--nnz", type=int, default=3)
    ap.add_argument("--gen-warps", type=int, default=0, dest="gen_warps", help="16 / 24 / 32 warps of the multi-step generate kernel")
    args = ap.parse_args()
    import torch
    import ecolab_b200 as E
    from helpers import SCALARS, synth
    peak = 6559.4
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = float(json.load(open(pk))["hbm_gbs"])
    m = E.SpatialModel(0, None)
    p = synth(args.nsp, args.nnz, 10, mig=0.1, mut=1e-5)
    for k, v in SCALARS.items():
        getattr(m, k)(v)
    for k, v in p.items():
        if k.startswith("interaction."):
            getattr(m.interaction, k.split(".")[1])(v)
        else:
            getattr(m, k)(v)
    nx = ny = args.n
    m.setGrid(nx, ny)
    m.seed(10)
    m.fill_density_random(0, 199)
    elems = nx * ny * args.nsp
    out = {"workload": f"spatial {nx}x{ny} cells x {args.nsp} species, {args.nnz} nnz/row", "hbm_peak_gbs": peak, "passes": {}}

    def timed(name, fn, bytes_per_elem, reps=3, note=""):
        best = None
        res = None
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            res = fn()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        gbs = bytes_per_elem * elems / best / 1e9
        out["passes"][name] = {"ms": best * 1e3, "algorithmic_bytes_per_element": bytes_per_elem, "achieved_gbs": gbs,
                               "frac_of_measured_hbm_peak": gbs / peak, "note": note}
        return res

    m.step(1)                                   # warm: second density buffer, beta upload
    timed("generate(1)", lambda: m.generate(1), 8)
    timed("migrate()", lambda: m.migrate(), 8)
    timed("generate(100) [per sub-step]", lambda: m.generate(100), 8 / 100.0, reps=1,
          note="one launch, 100 Jacobi steps per tile in shared memory; ms is for all 100")
    out["passes"]["generate(100) [per sub-step]"]["ms_per_substep"] = out["passes"]["generate(100) [per sub-step]"]["ms"] / 100
    timed("nsp() per cell", lambda: m.nsp(), 4, note="includes the D2H of one uint32 per cell")
    timed("species_totals()", lambda: m.species_totals(), 4)
    timed("alive mask (condense, nothing extinct)", lambda: m.alive_mask(), 4)
    # condense that really compacts: extinguish every 16th species everywhere
    nsp0 = int(m._scalar("nsp"))
    mask = np.ones(nsp0, dtype=np.uint8)
    mask[::16] = 0
    t = timed("compaction (condense_apply, 1/16 of the species removed)", lambda: m.condense_apply(mask) if int(m._scalar("nsp")) == nsp0 else None, 8, reps=1)
    del t
    # mutate: speciation scan over all cells + parameter draw + append
    m.mutation(np.full(int(m._scalar("nsp")), 1e-9)); m.mut_max(1e-9)   # ~50 speciation events in 10^9 elements
    before = int(m._scalar("nsp"))
    timed("mutate() (scan + host parameter draw + append)", lambda: m.mutate(), 4, reps=1)
    out["mutants_appended"] = int(m._scalar("nsp")) - before
    out["launches"] = m.launches()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
