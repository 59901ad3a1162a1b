#!/usr/bin/env python
"""Key metrics per kernel from an .ncu-rep (`ncu -i REP --page raw --csv`).
python profiles/ncu_key.py REP [more metric substrings]"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum ", "dram__bytes_write.sum ",
    "lts__t_bytes.sum ", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread ", "launch__block_size", "launch__grid_size",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "smsp__issue_active.avg.pct", "smsp__inst_executed.sum ",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum ",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ",
    "smsp__pcsamp_warps_issue_stalled", "lts__t_sector_hit_rate.pct",
]


def main():
    rep = sys.argv[1]
    extra = sys.argv[2:]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"],
                         capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_col = hdr.index("Kernel Name")
    for r in data:
        print("==", r[name_col][:80])
        stalls = []
        for i, h in enumerate(hdr):
            hh = h + " "
            if any(hh.startswith(k) for k in KEYS) or any(k in h for k in extra):
                if "pcsamp_warps_issue_stalled" in h:
                    try:
                        stalls.append((float(r[i].replace(",", "")), h.replace("smsp__pcsamp_warps_issue_stalled_", "")))
                    except ValueError:
                        pass
                    continue
                print(f"   {h:80s} {r[i]:>18s} {units[i]}")
        tot = sum(s for s, h in stalls if not h.endswith("not_issued")) or 1
        stalls = [(s, h) for s, h in stalls if not h.endswith("not_issued")]
        print("   stalls:", ", ".join(f"{h}={s/tot:.0%}" for s, h in sorted(stalls, reverse=True)[:7]))


if __name__ == "__main__":
    main()
