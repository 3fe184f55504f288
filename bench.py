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
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU leg (oracle): the reference's algorithm on the host cores.  Only place bench.py touches oracle/.
# ------------------------------------------------------------------------------------------------
def cpu_leg(nsp, nnz, steps, warmup, budget_s=20.0):
    from oracle import oracle as O
    from helpers import SCALARS, load_oracle
    O.build()
    threads = O.max_threads()
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
    for knob, val in (("gen_stages", args.gen_stages), ("gen_nobsmem", args.gen_nobsmem), ("mig_variant", args.mig_variant),
                      ("mig_stages", args.mig_stages), ("gen_dual", args.gen_dual)):
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
    moved = 0
    for _ in range(args.steps):
        moved = m.step(1)
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
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
    # DRAM traffic of that kernel from the committed ncu --set full capture (profiles/traffic.json: bytes per update
    # measured at C3, 1024 x 1024 x 1024, same kernel and nnz/row), scaled to this launch's updates
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        ent = json.load(open(tp)).get(f"{name}:nnz{args.nnz}")
        if ent and args.nsp == ent.get("nsp"):
            traffic = ent["dram_bytes_per_update"] * local_updates
            traffic_src = ent["source"]
    kname = {"generate": "generate_pipe_kernel", "migrate": "migrate_kernel"}.get(name, name)
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_ms": k_ms / max(k_n, 1),
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in kt.items() if v[1]},
                "step_frac_at_16B": (16.0 * updates_per_step / world * args.steps / (ms * 1e-3) / 1e9) / peak}

    del m   # frees the device buffers before the next leg allocates its own
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
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def e2e_leg(E, args, comm, local, stream, world, rank):
    """set_density_all(host) -> step(1) -> get_density_all(host): the call sequence of a host-resident user
    (what the reference's Python getters/setters do), pinned host buffers, copies inside the timed region."""
    import torch
    import torch.distributed as dist
    nx = ny = args.e2e_n
    m = build_model(E, args, nx, ny, comm, local, stream)
    nloc = (m._y1 - m._y0) * nx
    host_in = torch.randint(0, 200, (nloc, args.nsp), dtype=torch.int32).pin_memory()
    host_out = torch.empty((nloc, args.nsp), dtype=torch.int32).pin_memory()
    a_in, a_out = host_in.numpy(), host_out.numpy()
    import ctypes as C

    def one():
        if world == 1 and not args.e2e_unpipelined:
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
    steps = max(2, min(args.steps, 5))
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
    assert int(a_out.sum()) > 0
    val = nx * ny * args.nsp * steps / (ms * 1e-3)
    del m
    return {"value": val, "unit": UNIT, "h2d_bytes_per_step": int(nx * ny * args.nsp * 4), "d2h_bytes_per_step": int(nx * ny * args.nsp * 4 + 8),
            "workload": f"spatial {nx}x{ny} cells x {args.nsp} species through host buffers (bounded: pinned host copy of the full "
                        f"grid per step)", "steps": steps, "ms_per_step": ms / steps,
            "call": "eco_step_host (row-banded H2D/kernel/D2H pipeline)" if world == 1 and not args.e2e_unpipelined
                    else "eco_set_density_all + eco_step + eco_get_density_all"}


def run_dense(args):
    """BASELINE.json config 5: 256 x 256 cells x 8192 species, beta fully dense -- generate on the tensor cores
    (tcgen05 kind::i8 slices, ecolab_b200/csrc/k_dense.cuh) + migrate, and the FP64 sparse kernel on the same model."""
    import torch
    import ecolab_b200 as E
    import ctypes as C
    from helpers import SCALARS
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    nx, ny, nsp = args.nx or 256, args.ny or 256, args.nsp if args.nsp != 1024 else 8192
    rng = np.random.default_rng(10)
    m = E.SpatialModel(0, None)
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
        m._scalar("dense_prof0")     # enables the in-kernel cycle counters
    for _ in range(args.warmup):
        m.step(1)
    assert m._scalar("dense_active") == 1
    m.set_tuning("timing", 1)
    torch.cuda.synchronize()
    l0 = m.launches()
    sampler = ClockSampler(0); sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        m.step(1)
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
    peak = 2.0 * bf16      # int8 tensor rate = 2 x bf16 (nominal 4.5 vs 2.25 POP/s); measured bf16 sustained x 2
    launches_per_step = kt["dense_mma"][1] / args.steps
    achieved = 2.0 * macs / (mma_ms * launches_per_step * 1e-3) / 1e12
    roofline = {"bound": "tensor", "kernel": "generate_dense_kernel", "achieved": achieved, "peak": peak, "unit": "TOP/s (int8)",
                "frac": achieved / peak, "traffic": None,
                "peak_source": "2 x MEASURED_PEAKS.json bf16_tflops_sustained (int8 = 2 x bf16 rate; of measured)" if pk else "fallback",
                "executed_int8_macs_per_step": macs, "beta_slices": int(args.dense_slices or 3), "density_slices": q,
                "algorithmic_tflops": 2.0 * nsp * updates / (mma_ms * launches_per_step * 1e-3) / 1e12,
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in kt.items() if v[1]}}
    # the FP64 sparse kernel on the same dense model (config 5: "tcgen05 contraction path vs sparse path")
    sparse_ms = None
    if not args.no_cpu:
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
        names = ("mma_wait_tmem_empty", "mma_wait_smem_full", "mma_issue", "tiles", "epi_wait_tmem_full", "epi_work")
        roofline["cta0_cycles_last_launch"] = {n: int(m._scalar(f"dense_prof{i}")) for i, n in enumerate(names)}
    line = {"metric": METRIC, "value": updates * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "s8 x u8 -> s32 slices, f64 epilogue -> i32", "data": "synthetic",
            "config": {"workload": f"dense-interaction stress: spatial {nx}x{ny} cells x {nsp} species, beta fully dense, generate(1)+migrate() per step",
                       "l2": "inputs (%.1f GiB) larger than L2" % (updates * 4 / 2**30)},
            "clocks": clocks, "gpu_launches": int(launches), "roofline": roofline,
            "sparse_fp64_generate_ms": sparse_ms, "dense_generate_ms": (kt["dense_mma"][0] + kt["dense_pack"][0]) / args.steps}
    print(json.dumps(line))


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
    ap.add_argument("--e2e-n", type=int, default=1024, dest="e2e_n")
    ap.add_argument("--mig-stages", type=int, default=0, dest="mig_stages")
    ap.add_argument("--gen-stages", type=int, default=0, dest="gen_stages")
    ap.add_argument("--gen-nobsmem", type=int, default=0, dest="gen_nobsmem")
    ap.add_argument("--mig-variant", type=int, default=0, dest="mig_variant")
    ap.add_argument("--gen-dual", type=int, default=0, dest="gen_dual")
    ap.add_argument("--e2e-unpipelined", action="store_true", dest="e2e_unpipelined")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--dense", action="store_true", help="config 5: 256x256x8192, dense beta, tensor-core generate")
    ap.add_argument("--dense-slices", type=int, default=0, dest="dense_slices")
    ap.add_argument("--dense-prof", action="store_true", dest="dense_prof")
    ap.add_argument("--dense-dmax", type=int, default=199, dest="dense_dmax")
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
