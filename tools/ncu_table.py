#!/usr/bin/env python
"""Raw-page extract of an .ncu-rep (or of its exported raw CSV) as a markdown table:
    python tools/ncu_table.py file.ncu-rep|file_raw.csv > out.md"""
import csv, io, subprocess, sys
COLS = ["Kernel Name", "Grid Size", "gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__waves_per_multiprocessor"]
if sys.argv[1].endswith(".csv"):
    out = open(sys.argv[1]).read()           # a raw page already exported on the GPU box
else:
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = [(c, hdr.index(c)) for c in COLS if c in hdr]
print("| " + " | ".join(f"{c} [{units[i]}]" for c, i in idx) + " |")
print("|" + "---|" * len(idx))
for r in data:
    print("| " + " | ".join((r[i][:40] if c == "Kernel Name" else r[i]) for c, i in idx) + " |")
