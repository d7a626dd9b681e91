#!/usr/bin/env python
"""Key metrics per captured launch of one or more `ncu --set full` reports (reads them with
`ncu -i X.ncu-rep --page raw --csv`), in the form committed under profiles/.

    python tools/ncu_summary.py gpurun_out/r1j_norm.ncu-rep [more.ncu-rep ...] > profiles/r1j_norm_summary.txt
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "smsp__cycles_active.avg",
    "lts__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem",
]


def main(paths):
    for path in paths:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
        head, units, body = rows[hi], rows[hi + 1], rows[hi + 2:]
        ki = head.index("Kernel Name")
        print(f"== {path}")
        for r in body:
            if len(r) != len(head):
                continue
            print(f"-- {r[ki][:150]}")
            for k in KEYS:
                if k in head:
                    j = head.index(k)
                    print(f"   {k:82s} {r[j]:>18s} {units[j]}")
            for j, h in enumerate(head):   # tensor-pipe metrics are named differently across ncu versions: list them all
                if "pipe_tensor" in h and h not in KEYS and r[j] not in ("", "0", "n/a"):
                    print(f"   {h:82s} {r[j]:>18s} {units[j]}")


if __name__ == "__main__":
    main(sys.argv[1:])
