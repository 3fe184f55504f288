"""Runs ONE of the reference's driver scripts (tests/ref_scripts/, verbatim) through ecolab_b200.compat in this process --
plain `python` for one GPU, under torchrun for a row-slab partition -- stops it after a budget of generate() calls and
writes a fingerprint of the final model state (rank 0).  tests/test_gpu_ref_scripts.py compares the fingerprints of the
1-rank and the 2-rank run: the Philox stream is keyed on global cell ids and ModelData::mutate draws on rank 0 only, so
they must be identical.

    python tests/ref_script_worker.py <script.py> <budget> <out.json> [script args...]
"""
import hashlib
import json
import os
import runpy
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class StopScript(BaseException):
    pass


def main():
    script, budget, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    import ecolab_b200.compat  # noqa: F401
    import ecolab_model
    calls = {"n": 0}

    def budgeted(obj):
        real = obj.generate

        def generate(niter=1):
            if calls["n"] >= budget:
                raise StopScript()
            calls["n"] += 1
            return real(niter)

        obj._obj().generate = generate     # instance attribute shadows the method; everything else is untouched

    budgeted(ecolab_model.spatial_ecolab)
    budgeted(ecolab_model.panmictic_ecolab)
    os.environ.setdefault("ECOLAB_STEPS", "3")
    # the scripts fill their parameter lists from Python's process-global random(), unseeded: give every rank of every run
    # the same stream, or the 1-rank and the 2-rank run (and the two ranks of one run) start from different models
    import random
    random.seed(20251105)
    sys.argv = [script] + sys.argv[4:]
    try:
        runpy.run_path(script, run_name="__main__")
    except StopScript:
        pass
    spatial = "panmictic" not in open(script).read().split("\n", 3)[0] and "panmictic" not in open(script).read()[:400]
    m = (ecolab_model.spatial_ecolab if spatial else ecolab_model.panmictic_ecolab)._obj()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if spatial:
        loc = m.density_array()
        if world > 1:
            import torch
            import torch.distributed as dist
            parts = [None] * world
            dist.all_gather_object(parts, loc)
            dens = np.concatenate(parts)
            ns_all = m.nsp()
        else:
            dens, ns_all = loc, m.nsp()
    else:
        dens, ns_all = np.array([m.density()], dtype=np.int32), [m.nsp()]
    h = lambda a: hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()   # noqa: E731
    fp = {"generates": calls["n"], "tstep": int(m.tstep()), "nsp": len(m.species), "species": h(np.array(m.species(), dtype=np.int64)),
          "species_unique": len(set(m.species())) == len(m.species), "density": h(dens.astype(np.int64)), "individuals": int(dens.sum()), "nsp_per_cell": list(map(int, ns_all)),
          "repro_rate": h(np.array(m.repro_rate())), "val": h(np.array(m.interaction.val())), "world": world,
          "halo": "peer" if spatial and world > 1 and int(m._scalar("peer_connected")) else "-"}
    if rank == 0:
        json.dump(fp, open(out, "w"))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
