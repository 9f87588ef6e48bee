#!/usr/bin/env python3
"""Summarise an ``ncu --set full`` report for ``profiles/``: per kernel launch the duration, DRAM bytes, tensor-pipe and
issue utilisation, shared-memory wavefronts by source and the top warp stalls.

  python tools/ncu_summary.py gpurun_out/r01_v6_full.ncu-rep > profiles/r01_ncu_v6_summary.txt
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__cluster_size",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__inst_executed.avg.per_cycle_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.avg.per_second",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    print(f"# {rep}: {len(data)} launches (ncu --set full --clock-control none; per-launch values, cold caches, serialised)")
    for r in data:
        print(f"\nKernel: {r[col['Kernel Name']]}  id {r[col['ID']]}")
        for k in KEYS:
            hit = [h for h in hdr if h.endswith(k)]
            if hit:
                print(f"  {k} = {r[col[hit[0]]]} {units[col[hit[0]]]}")
        stalls = []
        for h in hdr:
            if "smsp__average_warps_issue_stalled_" in h and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
                try:
                    stalls.append((float(r[col[h]]), h.split("stalled_")[1].split("_per_issue")[0]))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        print("  top stalls (warps per issue): " + ", ".join(f"{n} {v:.2f}" for v, n in stalls[:6]))


if __name__ == "__main__":
    main()
