#!/usr/bin/env python
"""ncu -i <rep> --page raw --csv  ->  one line per launch with the counters the roofline uses."""
import csv
import subprocess
import sys

COLS = [
    ("grid", "launch__grid_size"), ("regs", "launch__registers_per_thread"),
    ("smemKB", "launch__shared_mem_per_block_dynamic"), ("waves", "launch__waves_per_multiprocessor"),
    ("us", "gpu__time_duration.sum"), ("rdMB", "dram__bytes_read.sum"), ("wrMB", "dram__bytes_write.sum"),
    ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("sm%", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("tensor%", "sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed"),
    ("tf32%", "sm__ops_path_tensor_op_utchmma_src_tf32_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed"),
    ("l2hit%", "lts__t_sector_hit_rate.pct"), ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("ipc", "sm__inst_executed.avg.per_cycle_elapsed"),
]


def tonum(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    print("# units: " + ", ".join(f"{n}={units[ix[m]]}" for n, m in COLS if m in ix))
    print(f"{'#':>3s} {'kernel':44s} " + " ".join(f"{n:>8s}" for n, _ in COLS))
    for n, r in enumerate(rows[2:]):
        name = r[ix["Kernel Name"]].replace("void ", "")[:44]
        vals = []
        for _, m in COLS:
            v = tonum(r[ix[m]]) if m in ix else None
            vals.append(f"{v:8.2f}" if v is not None else f"{'-':>8s}")
        print(f"{n:3d} {name:44s} " + " ".join(vals))


if __name__ == "__main__":
    main(sys.argv[1])
