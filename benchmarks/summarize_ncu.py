#!/usr/bin/env python
"""Condense an `ncu --set full` report into the CSV kept under profiles/.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv
    python benchmarks/summarize_ncu.py raw.csv profiles/ncu_<name>_raw_summary.csv
"""
import csv
import sys

KEEP = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "gpu__time_duration.sum", "sm__cycles_elapsed.avg",
    "sm__cycles_elapsed.avg.per_second", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
]


def main(src, dst):
    rows = list(csv.reader(open(src)))
    head, units = rows[0], rows[1]
    cols = [i for i, h in enumerate(head) if h in KEEP]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([head[i] for i in cols])
        w.writerow([units[i] for i in cols])
        for r in rows[2:]:
            w.writerow([r[i] for i in cols])
    print(f"{dst}: {len(rows) - 2} kernels x {len(cols)} metrics")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
