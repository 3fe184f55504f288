#!/usr/bin/env python
"""bench.py -- cell-species updates/s of one generate(1)+migrate() step (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]
                  [--nx X --ny Y --nsp S --nnz R] [--no-cpu] [--no-e2e]

One "step" = SpatialModel::generate(1) followed by SpatialModel::migrate() over the whole grid
(models/ecolab_model.cc:607-614, :616-703).  One "update" = one (cell, species) density advanced through that
step.  Workload: the grid the metric is quoted on, 4096 x 4096 cells x 1024 species (64 GiB of int32
densities) when it fits the device(s), else the largest square power-of-two grid that does; beta has
`--nnz` off-diagonal entries per row (default 3 = the reference scripts' random_interaction(3,0)).
For N > 1 the SAME grid is partitioned into row slabs (strong scaling), halo rows over NCCL.

Legs of the JSON line:
  value     : device-resident densities, K steps between CUDA events on the engine's stream (max over ranks)
  e2e       : the same step through the host-buffer API (eco_set_density_all -> eco_step -> eco_get_density_all),
              H2D + D2H of the full density inside the timed region, on a bounded grid (see e2e.workload)
  roofline  : dominant kernel's algorithmic bytes / its CUDA-event duration vs MEASURED_PEAKS.json hbm_gbs
  cpu_baseline : the CPU oracle (restatement of the reference's OpenMP path; the reference cannot be compiled
              here) on all host threads over a bounded sample of the same workload
--impl reference times ONLY that CPU leg and prints it as the reference arm.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

METRIC = "cell-species updates/sec (generate+migrate)"
UNIT = "updates/s"


def synth_params(nsp, nnz, seed=10):
    """SURVEY 8(d) synthetic inputs: r ~ U(-0.1,0.1), diag -1e-3, val ~ U(-1e-5,1e-5), nnz distinct columns/row."""
    from helpers import synth
    return synth(nsp, nnz, seed, mig=0.1)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        pw = [float(r[2]) for r in self.rows if len(r) >= 7 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "sm_min_mhz": min(sm) if sm else None,
                "power_w": float(np.median(pw)) if pw else None}


# ------------------------------------------------------------------------------------------------
# CPU leg (oracle): the reference's algorithm on the host cores.  Only place bench.py touches oracle/.
# ------------------------------------------------------------------------------------------------
def host_threads():
    """Every host thread this process may run on.  torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm is rank 0
    alone on the box, so it takes the affinity mask, not that variable."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_leg(nsp, nnz, steps, warmup, budget_s=20.0):
    threads = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(threads)     # before libgomp initialises
    from oracle import oracle as O
    from helpers import SCALARS, load_oracle
    O.build()
    O.set_threads(threads)
    p = synth_params(nsp, nnz)
    # bounded sample of the same workload: a 256 x H strip of the grid, H sized so the whole leg takes ~budget_s
    nx, ny = 256, 8
    d = np.full((nx * ny, nsp), 100, dtype=np.int32)
    o = load_oracle(O, p, d, (nx, ny), seed=10)
    t0 = time.perf_counter(); o.generate(1); o.migrate(); probe = time.perf_counter() - t0
    rate = nx * ny * nsp / probe
    target_updates = rate * budget_s / max(steps + warmup, 1)
    ny = int(max(8, min(4096, target_updates / (nx * nsp))))
    d = np.full((nx * ny, nsp), 100, dtype=np.int32)
    o = load_oracle(O, p, d, (nx, ny), seed=10)
    o.fill_density_random(0, 199)    # the same synthetic state as the GPU arm
    for _ in range(warmup):
        o.generate(1); o.migrate()
    t0 = time.perf_counter()
    for _ in range(steps):
        o.generate(1); o.migrate()
    dt = time.perf_counter() - t0
    val = nx * ny * nsp * steps / dt
    return {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{nx}x{ny} cells x {nsp} species, {nnz} nnz/row, {steps} steps of generate(1)+migrate() "
                      f"(oracle restatement of the reference OpenMP path, g++ -O3 -fopenmp -ffp-contract=off)"}, dt / steps * 1e3


def workload_name(nx, ny, nsp):
    return f"spatial {nx}x{ny} cells x {nsp} species, generate(1)+migrate() per step"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, ms = cpu_leg(args.nsp, args.nnz, args.steps, args.warmup, budget_s=60.0)
    nx, ny = (args.nx, args.ny) if args.nx and args.ny else (4096, 4096)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64->i32", "data": "synthetic",
            "config": {"workload": workload_name(nx, ny, args.nsp), "nnz_per_row": args.nnz, "sample": cb["sample"],
                       "note": "reference CPU algorithm on host cores, each step a bounded strip of the workload; the reference "
                               "build itself needs the absent classdesc submodule, so this is the line-by-line oracle restatement"},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
# native arm
# ------------------------------------------------------------------------------------------------
def pick_grid(args, world, free_bytes):
    if args.nx and args.ny:
        return args.nx, args.ny
    n = 4096
    while n > 64:
        need = 2 * (n * (n // world + 2)) * args.nsp * 4 + (2 << 30)   # two density buffers (migrate is out of place) + slack
        if need <= free_bytes * 0.92:
            break
        n //= 2
    return n, n


def build_model(E, args, nx, ny, comm, device, stream):
    import ctypes as C
    from helpers import SCALARS
    m = E.SpatialModel(device, comm)
    m.set_stream(stream)
    p = synth_params(args.nsp, args.nnz)
    for k, v in SCALARS.items():
        getattr(m, k)(v)
    for k, v in p.items():
        if k.startswith("interaction."):
            getattr(m.interaction, k.split(".")[1])(v)
        else:
            getattr(m, k)(v)
    m.setGrid(nx, ny)
    m.seed(10)
    for knob, val in (("gen_stages", args.gen_stages), ("gen_nobsmem", args.gen_nobsmem), ("mig_rows", args.mig_rows), ("gen_dual", args.gen_dual)):
        if val:
            m.set_tuning(knob, val)
    m.fill_density_random(0, 199)   # mean 100, with gradients between neighbours so migrate really moves population
    return m


def run_native(args):
    import torch
    import torch.distributed as dist
    import ecolab_b200 as E
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    comm = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        comm = E.Comm()
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    free, total = torch.cuda.mem_get_info()
    nx, ny = pick_grid(args, world, free)
    m = build_model(E, args, nx, ny, comm, local, stream)
    updates_per_step = nx * ny * args.nsp

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        m.step(1)
    m.set_tuning("timing", 1)
    barrier()
    l0 = m.launches()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        m.step(1, wait=False)       # no host round trip per step: migrate's total is read once, after the loop
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    moved = m.last_migration()
    ms = e0.elapsed_time(e1)
    launches = m.launches() - l0
    kt = m.kernel_times()
    m.set_tuning("timing", 0)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        lt = torch.tensor([launches], device="cuda", dtype=torch.int64)
        dist.all_reduce(lt, op=dist.ReduceOp.SUM)
        launches = int(lt.item())
    value = updates_per_step * args.steps / (ms * 1e-3)

    # ---- roofline of the dominant kernel (CUDA events around every launch, on the engine's stream) ----
    peak, peak_src = peaks()
    local_updates = (m._y1 - m._y0) * nx * args.nsp
    name = max(kt, key=lambda k: kt[k][0])
    k_ms, k_n = kt[name]
    bytes_per_launch = 8.0 * local_updates            # 4 B read + 4 B written per cell-species update
    achieved = bytes_per_launch / (k_ms / max(k_n, 1) * 1e-3) / 1e9 if k_ms > 0 else 0.0
    # DRAM traffic of that kernel from the committed ncu captures (profiles/traffic.json)
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        # the capture of a launch of this grid width whose row count is closest to this rank's (C4 whole, a 512-row slab of it, C3)
        cand = [e for e in json.load(open(tp)).get("entries", []) if e["kernel"] == name and e["nnz"] == args.nnz and e["nsp"] == args.nsp]
        rows = m._y1 - m._y0
        exact = sorted((e for e in cand if e["grid"][0] == nx), key=lambda e: abs(e["grid"][1] - rows))
        if exact:
            traffic = exact[0]["dram_bytes_per_update"] * local_updates
            traffic_src = exact[0]["source"]
    kname = {"generate": "generate_pipe_kernel", "migrate": "migrate_kernel"}.get(name, name)
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_ms": k_ms / max(k_n, 1),
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in kt.items() if v[1]},
                "step_frac_at_16B": (16.0 * updates_per_step / world * args.steps / (ms * 1e-3) / 1e9) / peak}

    # ---- verify (outside the timed region): partition-independent fingerprints of the state + band parity vs the oracle ----
    verify = None
    if not args.no_verify:
        verify = verify_leg(m, args, nx, ny, world, rank)

    m.close()   # frees the device buffers before the next leg allocates its own
    del m
    # ---- e2e: host buffers through the C ABI, H2D + D2H inside the timed region, bounded grid ----
    e2e = None
    if not args.no_e2e:
        e2e = e2e_leg(E, args, comm, local, stream, world, rank)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu, _ = cpu_leg(args.nsp, args.nnz, 3, 1, budget_s=15.0)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64->i32", "data": "synthetic",
                "config": {"workload": workload_name(nx, ny, args.nsp),
                           "nnz_per_row": args.nnz, "partition": f"{world} row slab(s)", "l2": "inputs (%.1f GiB) larger than L2" % (updates_per_step * 4 / 2**30),
                           "rng": "Philox4x32-10 keyed (seed, cell, slot, tstep)", "migrated_last_step": int(moved)},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
                "verify": verify}
        extra = {}
        if world == 1 and not args.no_extra:
            extra = extra_legs(E, args, local, stream)
        line["extra"] = extra
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def verify_leg(m, args, nx, ny, world, rank):
    """Two more steps, each checked on a 4-row band against the oracle fed with the GPU's own pre-step rows (top of the
    grid; the middle, where every even partition has a seam), then fingerprints of the state that do not depend on the
    partition: the driver can compare the lines of N = 1/2/4/8.  The oracle is the checker here, never the thing timed."""
    import hashlib
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    from helpers import SCALARS, band_step_equals_oracle, collect_rows
    if rank == 0:
        O.build()
    p = synth_params(args.nsp, args.nnz)
    bands = [0, max(0, ny // 2 - 2)]
    nb = min(4, ny)
    ok = []
    for y in bands:
        eq, _ = band_step_equals_oracle(O, m, p, SCALARS, 10, y, nb)
        ok.append(eq)
    tot = torch.from_numpy(m.species_totals()).cuda()
    if world > 1:
        dist.all_reduce(tot)
    rows = [collect_rows(m, y, nb) for y in bands]
    if rank != 0:
        return None
    h = lambda a: hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]   # noqa: E731
    return {"tstep": int(m.tstep()), "band_rows": [[y, y + nb] for y in bands], "band_equal_oracle": bool(all(ok)),
            "band_sha1": [h(r) for r in rows], "species_totals_sha1": h(tot.cpu().numpy()),
            "individuals": int(tot.sum().item()),
            "note": "after warmup+steps+2 steps; bands checked against the CPU oracle run on the GPU's own pre-step rows"}


def extra_legs(E, args, local, stream):
    """SURVEY 8(d) legs beside the headline (N = 1 only, each bounded to a few seconds): the nnz/row sweep of generate on
    C3, the scripts' real cadence generate(100)+mutate()+migrate()+condense() on C2 with the scripts' parameters (this is
    the leg that times the host-serial ModelData::mutate/condense stage), and the dense config 5 on the tensor cores."""
    import copy
    import torch
    out = {}
    sweep = {}
    for nnz in (3, 16, 51):
        a = copy.copy(args); a.nnz = nnz
        m = build_model(E, a, 1024, 1024, None, local, stream)
        for _ in range(3):
            m.step(1, wait=False)
        m.set_tuning("timing", 1)
        torch.cuda.synchronize()
        for _ in range(5):
            m.step(1, wait=False)
        torch.cuda.synchronize()
        kt = m.kernel_times()
        m.set_tuning("timing", 0)
        g, mg = kt["generate"][0] / 5, kt["migrate"][0] / 5
        sweep[str(nnz)] = {"generate_ms": g, "migrate_ms": mg, "updates_per_s": 2 ** 30 / ((g + mg) * 1e-3),
                           "generate_hbm_frac": 8.0 * 2 ** 30 / (g * 1e-3) / 1e9 / peaks()[0]}
        m.close(); del m
    out["nnz_sweep_c3"] = {"workload": "spatial 1024x1024 cells x 1024 species, generate(1)+migrate(), 5 steps", "by_nnz_per_row": sweep}
    # the scripts' step on a big grid: generate(100) then ONE migrate (models/steady_migration.py:63-71) -- the 100 sub-steps of a
    # tile run on chip (generate_multi_kernel), so HBM sees 8/100 B per update-step
    a = copy.copy(args); a.nnz = 3
    m = build_model(E, a, 1024, 1024, None, local, stream)
    m.step(100, wait=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(2):
        m.step(100, wait=False)
    e1.record(stream)
    torch.cuda.synchronize()
    ms100 = e0.elapsed_time(e1) / 2
    out["generate100_migrate_c3"] = {"workload": "spatial 1024x1024 cells x 1024 species, 3 nnz/row: generate(100)+migrate() per round, 2 rounds",
                                     "ms_per_round": ms100, "update_steps_per_s": 2 ** 30 * 100 / (ms100 * 1e-3)}
    m.close(); del m
    # the scripts' cadence (models/spatial_ecolab.py:66-78) at BASELINE.json config 2's shape
    a = copy.copy(args); a.nsp = 1000; a.nnz = 3
    rounds = 4
    for rep in range(2):   # the first pass is the warm-up: the species count grows through three tile widths, and CUDA loads
        m = build_model(E, a, 64, 64, None, local, stream)   # each kernel instantiation lazily on its first launch (~100 ms)
        m.mut_max(1e-5); m.mutation(np.full(1000, 1e-5))
        m.fill_density(100)                                   # the scripts' initial state (models/spatial_ecolab.py:45)
        E.lib().eco_array_urand_seed(10)
        walls, dev, nsps = [], [], []
        for r in range(rounds):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); m.generate(100); e1.record(stream)
            m.mutate(); m.migrate(); m.condense()
            torch.cuda.synchronize()
            walls.append(time.perf_counter() - t0); dev.append(e0.elapsed_time(e1) * 1e-3); nsps.append(len(m.species))
        if rep == 0:
            m.close(); del m
    upd = 64 * 64 * 1000 * 100
    out["cadence_c2"] = {"workload": "spatial 64x64 cells x 1000 species (3 nnz/row, mutation = mut_max = 1e-5, the scripts' parameters): "
                                     "generate(100)+mutate()+migrate()+condense() per round",
                         "rounds": rounds, "wall_ms_per_round": [w * 1e3 for w in walls], "generate100_ms_per_round": [d * 1e3 for d in dev],
                         "species_after_round": nsps,
                         "generation_updates_per_s": upd / float(np.median(walls)),
                         "non_generate_share": 1.0 - float(np.median(dev)) / float(np.median(walls))}
    m.close(); del m
    if not args.no_dense:
        try:
            out["dense_c5"] = dense_leg(E, args, local, stream, steps=3, warmup=2)
        except Exception as ex:   # noqa: BLE001
            out["dense_c5"] = {"unavailable": str(ex)[:200]}
    return out


def e2e_grid(args, world):
    """The grid of the e2e leg and whether the step runs IN PLACE (one pinned host buffer is both input and output: the
    banded pipeline writes band k-1 back after band k has been read).  The metric's own grid (C4) when one host copy of it,
    split over the ranks, stays within 64 GiB of pinned memory per rank and 40 % of the box's available RAM in total; else
    the largest square power-of-two grid whose separate input and output copies stay within 32 GiB per rank and 35 %."""
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:   # noqa: BLE001
        avail = 64 << 30
    if args.e2e_n:
        return args.e2e_n, False
    one = 4096 * 4096 * args.nsp * 4
    if one / world <= (64 << 30) and one <= 0.40 * avail and not args.e2e_unpipelined:
        return 4096, True
    n = 4096
    while n > 256:
        total = 2 * n * n * args.nsp * 4          # in + out
        if total / world <= (32 << 30) and total <= 0.35 * avail:
            break
        n //= 2
    return n, False


def e2e_leg(E, args, comm, local, stream, world, rank):
    """The step through the host-buffer API: pinned HOST densities in, pinned HOST densities out, H2D + kernels + D2H
    inside the timed region (what a host-resident user of the reference's getters/setters pays).  eco_step_host streams
    bands of grid rows; under torch.distributed every rank streams its own slab over its own PCIe link."""
    import torch
    import torch.distributed as dist
    n, inplace = e2e_grid(args, world)
    nx = ny = n
    m = build_model(E, args, nx, ny, comm, local, stream)
    nloc = (m._y1 - m._y0) * nx
    host_in = torch.empty((nloc, args.nsp), dtype=torch.int32, pin_memory=True)
    host_out = host_in if inplace else torch.empty((nloc, args.nsp), dtype=torch.int32, pin_memory=True)
    a_in, a_out = host_in.numpy(), host_out.numpy()
    block = np.random.default_rng(5 + rank).integers(0, 200, size=(min(nloc, 16384), args.nsp), dtype=np.int32)
    for i in range(0, nloc, len(block)):          # synthetic state, mean 100, tiled from one random block
        a_in[i:i + len(block)] = block[: min(len(block), nloc - i)]
    import ctypes as C

    def one():
        if not args.e2e_unpipelined:
            return m.step_host(a_in, a_out, 1)     # eco_step_host: banded H2D / kernels / D2H pipeline
        E._cabi.check(m._L.eco_set_density_all(m._h, C.c_void_p(a_in.ctypes.data), nloc, args.nsp))
        moved = m.step(1)
        E._cabi.check(m._L.eco_get_density_all(m._h, C.c_void_p(a_out.ctypes.data), nloc, args.nsp))
        return moved

    for _ in range(2):
        one()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    steps = 2 if inplace else max(2, min(args.steps, 4))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        one()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    assert int(a_out[:1024].sum()) > 0
    val = nx * ny * args.nsp * steps / (ms * 1e-3)
    m.close(); del m
    return {"value": val, "unit": UNIT, "h2d_bytes_per_step": int(nx * ny * args.nsp * 4), "d2h_bytes_per_step": int(nx * ny * args.nsp * 4 + 8 * world),
            "workload": f"spatial {nx}x{ny} cells x {args.nsp} species through pinned host buffers"
                        f"{' (in place: one buffer is input and output)' if inplace else ''}, {world} row slab(s), "
                        f"{nloc * args.nsp * 4 / 2**30:.1f} GiB per rank each way per step", "steps": steps, "ms_per_step": ms / steps,
            "call": "eco_step_host (row-banded H2D/kernel/D2H pipeline; partitioned: eco_step_host_begin + halo + eco_step_host)"
                    if not args.e2e_unpipelined else "eco_set_density_all + eco_step + eco_get_density_all"}


def dense_leg(E, args, local, stream, steps, warmup, sparse_too=False):
    """BASELINE.json config 5: 256 x 256 cells x 8192 species, beta fully dense -- generate on the tensor cores
    (tcgen05 kind::i8 slices, ecolab_b200/csrc/k_dense.cuh) + migrate; optionally the FP64 sparse kernel on the same model."""
    import torch
    from helpers import SCALARS
    nx, ny, nsp = args.dense_n, args.dense_n, args.dense_nsp
    rng = np.random.default_rng(10)
    m = E.SpatialModel(local, None)
    m.set_stream(stream)
    for k, v in SCALARS.items():
        getattr(m, k)(v)
    m.species(np.arange(nsp, dtype=np.int32))
    m.repro_rate(rng.uniform(-0.1, 0.1, nsp))
    m.interaction.diag(np.full(nsp, -1e-3))
    idx = np.arange(nsp, dtype=np.uint32)
    row = np.repeat(idx, nsp - 1)
    col = np.tile(idx, nsp).reshape(nsp, nsp)[~np.eye(nsp, dtype=bool)].astype(np.uint32)
    # 67 M entries: the plain array setter of the host mirror (the callable members also return the list they set)
    m._set_array("interaction.row", "u32", row); m._set_array("interaction.col", "u32", col)
    m._set_array("interaction.val", "f64", rng.uniform(-1e-5, 1e-5, row.size))
    del row, col
    m.mutation(np.full(nsp, 1e-5)); m.migration(np.full(nsp, 0.1)); m.create(np.zeros(nsp))
    m.setGrid(nx, ny)
    m.seed(10)
    m.set_tuning("gen_dense", 1)        # the tensor path is opt-in (not bit-exact against the FP64 summation order)
    if args.dense_slices:
        m.set_tuning("dense_slices", args.dense_slices)
    m.fill_density_random(0, args.dense_dmax)   # < 256: one u8 slice per density; larger: two
    updates = nx * ny * nsp
    if args.dense_prof:
        m._scalar("dense_prof0")     # per-stage cycle counters of the issuing thread (tools: they slow it down)
    else:
        m._scalar("dense_clock_mhz")   # two clock reads per launch: the SM clock the kernel really runs at
    for _ in range(warmup):
        m.step(1, wait=False)
    assert m._scalar("dense_active") == 1
    m.set_tuning("timing", 1)
    torch.cuda.synchronize()
    l0 = m.launches()
    sampler = ClockSampler(local); sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        m.step(1, wait=False)
    e1.record(stream)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    kt = m.kernel_times()
    m.set_tuning("timing", 0)
    launches = m.launches() - l0
    macs, q = m._scalar("dense_macs"), int(m._scalar("dense_q"))
    mma_ms = kt["dense_mma"][0] / max(kt["dense_mma"][1], 1)
    pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    bf16 = float(pk.get("bf16_tflops_sustained", 1400.0))
    launches_per_step = kt["dense_mma"][1] / steps
    achieved = 2.0 * macs / (mma_ms * launches_per_step * 1e-3) / 1e12
    alg = 2.0 * nsp * updates / (mma_ms * launches_per_step * 1e-3) / 1e12
    # No measured int8 tensor peak exists in MEASURED_PEAKS.json (the driver measures bf16 only), so the roofline is quoted
    # on the ALGORITHMIC rate -- 2*nsp flops per update, what a bf16 GEMM of the same shape would be credited -- against
    # the measured bf16 peak; the executed int8 rate (P*Q slices) is reported beside it without a fraction.
    roofline = {"bound": "tensor", "kernel": "generate_dense_kernel", "achieved": alg, "peak": bf16, "unit": "TFLOP/s",
                "frac": alg / bf16, "traffic": None,
                "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (of measured); achieved = algorithmic 2*nsp flops per update" if pk else "fallback",
                "executed_int8_tops": achieved, "executed_int8_macs_per_step": macs,
                "beta_slices": int(args.dense_slices or 3), "density_slices": q,
                "kernels_ms_per_step": {k: v[0] / steps for k, v in kt.items() if v[1]}}
    # the int8 ceiling measured with tools/ubench_i8peak.cu (bare kind::i8 MMAs from resident random operands, all SMs, under
    # the part's power cap), and the bound that actually holds the kernel: shared-memory bandwidth.  Per 128-byte
    # K chunk of a tile the bulk copies WRITE (P + Q) x 16 KB into shared memory and the MMAs READ their operands from it -- an
    # M128 x N x 32 B MMA reads 4 KB of beta and N/128 x 4 KB of densities -- at 128 B/clk/SM for both together.
    ipk = os.path.join(ROOT, "profiles", "int8_peak.json")
    if os.path.exists(ipk):
        ip = json.load(open(ipk))
        roofline["int8_peak_tops"] = ip["tops_random_operands_n128"]
        roofline["int8_frac"] = achieved / ip["tops_random_operands_n128"]
        roofline["int8_peak_source"] = ip["source"]
        mhz = m._scalar("dense_clock_mhz") or clocks.get("sm_mhz")
        if mhz and q:
            P = int(args.dense_slices or 3)
            stages = macs / (P * q) / (128.0 * 128 * 128)                      # (tile, K chunk) pairs of one step, all SMs
            reads = 4 * P * (4096 + q * 4096) if q <= 2 else 4 * P * q * 8192   # q <= 2: the density slices ride one N = 128q MMA
            writes = (P + q) * 16384
            floor_ms = stages * (reads + writes) / (ip["shared_memory_bytes_per_clk_per_sm"] * 148 * mhz * 1e3)
            roofline["shared_memory_bound"] = {"bytes_per_stage": {"operand_reads": reads, "bulk_copy_writes": writes},
                                               "bytes_per_clk_per_sm": ip["shared_memory_bytes_per_clk_per_sm"], "sm_mhz_in_kernel": mhz,
                                               "floor_ms": floor_ms, "frac": floor_ms / (mma_ms * launches_per_step),
                                               "tensor_pipe_ceiling_in_this_tile_shape": 4 * P * (64 if q == 1 else 128 if q == 2 else 64 * q) * 128.0 / (reads + writes)}
    sparse_ms = None
    if sparse_too:   # the FP64 sparse kernel on the same dense model (config 5: "tcgen05 contraction path vs sparse path")
        try:
            m.set_tuning("gen_dense", -1)
            m.generate(1)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream); m.generate(1); b.record(stream)
            torch.cuda.synchronize()
            sparse_ms = a.elapsed_time(b)
        except Exception as ex:   # noqa: BLE001
            sparse_ms = f"unavailable: {ex}"
    if args.dense_prof:
        names = ("mma_wait_tmem_empty", "mma_wait_smem_full", "mma_issue", "tiles", "epi_wait_tmem_full", "epi_work", "epi_drain", "mma_loop_ns")
        roofline["cta0_cycles_last_launch"] = {n: int(m._scalar(f"dense_prof{i}")) for i, n in enumerate(names)}
    m.close(); del m
    return {"metric": METRIC, "value": updates * steps / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": steps,
            "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "s8 x u8 -> s32 slices, f64 epilogue -> i32", "data": "synthetic",
            "config": {"workload": f"dense-interaction stress: spatial {nx}x{ny} cells x {nsp} species, beta fully dense, generate(1)+migrate() per step",
                       "l2": "inputs (%.1f GiB) larger than L2" % (updates * 4 / 2**30)},
            "clocks": clocks, "gpu_launches": int(launches), "roofline": roofline,
            "sparse_fp64_generate_ms": sparse_ms, "dense_generate_ms": (kt["dense_mma"][0] + kt["dense_pack"][0]) / steps}


def run_dense(args):
    import torch
    import ecolab_b200 as E
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    print(json.dumps(dense_leg(E, args, 0, stream, args.steps, args.warmup, sparse_too=not args.no_cpu)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--nx", type=int, default=0)
    ap.add_argument("--ny", type=int, default=0)
    ap.add_argument("--nsp", type=int, default=1024)
    ap.add_argument("--nnz", type=int, default=3)
    ap.add_argument("--e2e-n", type=int, default=0, dest="e2e_n", help="grid edge of the e2e leg (0 = the metric's grid where the host copy fits)")
    ap.add_argument("--mig-rows", type=int, default=0, dest="mig_rows")
    ap.add_argument("--gen-stages", type=int, default=0, dest="gen_stages")
    ap.add_argument("--gen-nobsmem", type=int, default=0, dest="gen_nobsmem")
    ap.add_argument("--gen-dual", type=int, default=0, dest="gen_dual")
    ap.add_argument("--e2e-unpipelined", action="store_true", dest="e2e_unpipelined")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-verify", action="store_true", dest="no_verify")
    ap.add_argument("--no-extra", action="store_true", dest="no_extra")
    ap.add_argument("--no-dense", action="store_true", dest="no_dense")
    ap.add_argument("--dense", action="store_true", help="config 5: 256x256x8192, dense beta, tensor-core generate")
    ap.add_argument("--dense-slices", type=int, default=0, dest="dense_slices")
    ap.add_argument("--dense-prof", action="store_true", dest="dense_prof")
    ap.add_argument("--dense-dmax", type=int, default=199, dest="dense_dmax")
    ap.add_argument("--dense-n", type=int, default=256, dest="dense_n")
    ap.add_argument("--dense-nsp", type=int, default=8192, dest="dense_nsp")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: native libraries that write to file descriptor 1 (NCCL prints its version
    # there) are diverted to stderr, Python's own stdout keeps the original descriptor
    sys.stdout.flush()
    sys.stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    elif args.dense:
        run_dense(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
