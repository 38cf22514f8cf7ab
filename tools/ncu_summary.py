#!/usr/bin/env python3
"""Summarise an .ncu-rep (captured on the GPU box with `ncu --set full --import-source on`) into
a small text file for profiles/: headline counters from the raw page, the stall-reason totals and
the hottest SASS instructions from the source page.

usage: python tools/ncu_summary.py gpurun_out/prof_X.ncu-rep [> profiles/rNN_X.txt]
"""
import csv
import io
import subprocess
import sys

RAW_KEYS = [
    "Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "launch__waves_per_multiprocessor", "sm__maximum_warps_per_active_cycle_pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__pcsamp_sample_buffer_full.sum",
]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True)
    return out.stdout


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    rows = list(csv.reader(io.StringIO(page(rep, "raw"))))
    hdr, units = rows[0], rows[1]
    print(f"# ncu summary of {rep}")
    for launch in rows[2:]:
        print("## launch")
        for i, h in enumerate(hdr):
            if h in RAW_KEYS:
                print(f"{h:78s} {launch[i]:>16s} {units[i]}")
        # stall breakdown (per-warp issue stall ratios)
        for i, h in enumerate(hdr):
            if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio") \
                    and "not_issued" not in h:
                try:
                    v = float(launch[i])
                except ValueError:
                    continue
                if v >= 0.05:
                    print(f"{h:78s} {v:16.3f}")
    src = list(csv.reader(io.StringIO(page(rep, "source"))))
    # first line: kernel name; second: header
    start = next(i for i, r in enumerate(src) if r and r[0] == "Address")
    h = src[start]
    col = {n: i for i, n in enumerate(h)}
    stall_cols = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
    body = [r for r in src[start + 1:] if len(r) == len(h)]
    tot = {n: 0 for n in stall_cols}
    samples = 0
    for r in body:
        samples += int(r[col["# Samples"]] or 0)
        for n in stall_cols:
            tot[n] += int(r[col[n]] or 0)
    print(f"## warp-stall samples by reason (total {samples}, {len(body)} SASS instructions)")
    for n, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        if v:
            print(f"{n:28s} {v:8d} {100.0 * v / max(1, samples):6.1f}%")
    print(f"## top {top} instructions by samples")
    ranked = sorted(range(len(body)), key=lambda i: -int(body[i][col["# Samples"]] or 0))[:top]
    for i in sorted(ranked):
        r = body[i]
        s = int(r[col["# Samples"]] or 0)
        why = sorted(((int(r[col[n]] or 0), n) for n in stall_cols), reverse=True)[:2]
        why = ", ".join(f"{n[6:]}={v}" for v, n in why if v)
        print(f"[{i:5d}] {s:7d} {100.0 * s / max(1, samples):5.1f}%  {r[col['Source']].strip():70s} {why}")


if __name__ == "__main__":
    main()
