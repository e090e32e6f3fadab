"""Summarise an `ncu --page raw --csv` export: python tools/ncu_extract.py report.raw.csv > profiles/xxx.txt"""
import csv
import sys

KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__cluster",
        "sm__pipe_tensor_cycles_active", "sm__inst_executed_pipe_tensor", "sm__throughput.avg.pct",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct",
        "lts__t_bytes.sum", "lts__throughput.avg.pct", "l1tex__throughput.avg.pct",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warp_issue_stalled", "sm__warps_active.avg.pct",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active"]

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, vals = rows[0], rows[1], rows[2]
print("kernel:", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
for i, h in enumerate(hdr):
    if any(k in h for k in KEYS) and ".per_second" not in h and not h.endswith((".min", ".max")) \
            and ".min." not in h and ".max.pct" not in h and ".sum.pct" not in h:
        print("%-110s %-14s %s" % (h, units[i], vals[i]))
