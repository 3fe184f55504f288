#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, on the CPU box): the metrics the roofline argument uses plus the
top stall sites per source line / SASS instruction.

  python tools/ncu_summary.py gpurun_out/x.ncu-rep [--sass N] > profiles/x.txt
"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_config_size", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "lts__t_bytes.sum", "l1tex__t_bytes.sum", "sm__cycles_elapsed.avg",
]


def ncu(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    nsass = int(sys.argv[sys.argv.index("--sass") + 1]) if "--sass" in sys.argv else 25
    rows = ncu(rep, "raw")
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("kernel:", r[hdr.index("Kernel Name")], " grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
        for h, u, v in zip(hdr, units, r):
            if h in WANT:
                print(f"  {h:75s} {v} {u}")
        stalls = [(float(v or 0), h) for h, v in zip(hdr, r) if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
        print("  stall reasons (warps stalled per issue-active cycle):")
        for v, h in sorted(stalls, reverse=True)[:8]:
            print(f"    {h.split('issue_stalled_')[1].split('_per_issue')[0]:28s} {v:.3f}")
    # per-instruction view
    src = ncu(rep, "source", ("--print-source", "sass"))
    hi = next((i for i, r in enumerate(src) if "Source" in r and "Address" in r), None)
    if hi is not None:
        h = src[hi]
        src = src[hi:]
        def col(name):
            for i, c in enumerate(h):
                if c.strip() == name:
                    return i
            return None
        ci, cs, cx = col("Source"), col("Warp Stall Sampling (All Samples)"), col("Instructions Executed")
        if ci is not None and cs is not None:
            tot = sum(float(r[cs] or 0) for r in src[1:] if len(r) > cs and r[cs].replace('.', '').isdigit())
            ranked = sorted(((float(r[cs] or 0), i, r) for i, r in enumerate(src[1:]) if len(r) > cs and r[cs].replace('.', '').isdigit()), reverse=True)
            print(f"top SASS stall sites (of {tot:.0f} samples):")
            for v, i, r in ranked[:nsass]:
                why = ""
                sc = [(float(r[j] or 0), h[j]) for j in range(len(h)) if h[j].startswith("stall_") and "Not Issued" not in h[j] and j < len(r)]
                if sc:
                    why = max(sc)[1]
                print(f"  {100 * v / max(tot, 1):5.1f}%  #{i:4d} exec={r[cx] if cx is not None else '?':>10s} {why:18s} {r[ci].strip()[:80]}")


if __name__ == "__main__":
    main()
