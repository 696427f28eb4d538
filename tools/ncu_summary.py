#!/usr/bin/env python
"""ncu -i <rep> --page raw --csv  ->  a compact per-launch table (profiles/*.txt)."""
import csv
import subprocess
import sys

WANT = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
    "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "lts__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def main(rep, extra=()):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    want = list(WANT) + [e for e in extra]
    for h in hdr:
        if ("tensor" in h or "umma" in h) and h not in want:
            want.append(h)
    idx = [hdr.index(w) for w in want if w in hdr]
    for n, r in enumerate(rows[2:]):
        print(f"--- launch {n}")
        for i in idx:
            print(f"{hdr[i]:84s} {r[i]:>18s} {units[i]}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
