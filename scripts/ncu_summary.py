#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): key raw metrics + top stall reasons per source line.

  python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [--source N]
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__cycles_active.avg", "sm__cycles_elapsed.avg", "sm__inst_executed.sum", "sm__inst_executed.sum.per_cycle_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.avg.per_cycle_active", "l1tex__data_bank_conflicts_pipe_lsu.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.pct", "smsp__warps_eligible.avg.per_cycle_active",
]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for row in rows[2:]:
        name = row[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        print(f"## {name[:100]}")
        d = {h: (row[i], units[i]) for i, h in enumerate(hdr)}
        for k in KEYS:
            if k in d:
                print(f"  {k:78s} {d[k][0]:>16s} {d[k][1]}")
        stalls = sorted(((float(v[0].replace(',', '')), k) for k, v in d.items()
                         if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and v[0]),
                        reverse=True)
        print("  -- warp stall reasons (warps stalled per issue-active cycle), top 8")
        for val, k in stalls[:8]:
            print(f"     {k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''):32s} {val:8.3f}")


def source(rep, top):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    while rows and "Source" not in rows[0]:
        rows = rows[1:]
    if not rows:
        return
    hdr = rows[0]
    def col(name):
        for i, h in enumerate(hdr):
            if h.strip() == name:
                return i
        return None
    c_src, c_samp, c_inst = col("Source"), col("# Samples"), col("Instructions Executed")
    if c_src is None or c_samp is None:
        print("  (no per-instruction sampling columns found)", hdr[:12])
        return
    items = []
    for r in rows[1:]:
        try:
            items.append((int(r[c_samp].replace(',', '') or 0), r[c_src].strip()[:110]))
        except (ValueError, IndexError):
            pass
    tot = sum(s for s, _ in items) or 1
    print(f"  -- top {top} instructions by stall samples (total {tot})")
    for s, src in sorted(items, reverse=True)[:top]:
        print(f"     {100.0 * s / tot:5.1f}%  {src}")


if __name__ == "__main__":
    rep = sys.argv[1]
    raw(rep)
    if "--source" in sys.argv:
        source(rep, int(sys.argv[sys.argv.index("--source") + 1]))
