#!/bin/bash
# ncu --set full of ONE kernel (regex $1, launch skip $2) of a command ($3...), raw page -> gpurun_out/ncu_one_raw.csv + key rows
K=$1; S=$2; shift 2
ncu --set full --clock-control none --import-source on -k regex:"$K" -s $S -c 1 -f -o gpurun_out/ncu_one "$@" > gpurun_out/ncu_one.log 2>&1
ncu -i gpurun_out/ncu_one.ncu-rep --page raw --csv > gpurun_out/ncu_one_raw.csv 2>/dev/null
python - <<'EOF2'
import csv
rows = list(csv.reader(open("gpurun_out/ncu_one_raw.csv")))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__inst_executed_pipe", "smsp__average_warp",
        "smsp__average_warps_issue_stalled", "dram__bytes", "dram__throughput.avg.pct", "l1tex__t_sector_hit_rate", "lts__t_sector_hit_rate",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct", "local", "l1tex__data_bank_conflicts"]
for r in data:
    for i, h in enumerate(hdr):
        if any(h.startswith(k) or (k in h and k in ("local",)) for k in keys):
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                v = r[i]
            if v not in (0, 0.0, "0"):
                print(f"{h} [{units[i]}] = {r[i]}")
EOF2
rm -f gpurun_out/ncu_one.ncu-rep
