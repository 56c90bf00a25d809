mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none -k regex:wgrad_tc_kernel -s 3 -c 1 -f -o gpurun_out/prof_wg python tools/wgrad_tune.py 32 > gpurun_out/prof_wg.log 2>&1; echo "rc=$?"
ncu -i gpurun_out/prof_wg.ncu-rep --page raw --csv > gpurun_out/prof_wg_raw.csv 2>/dev/null
ncu -i gpurun_out/prof_wg.ncu-rep --page details 2>/dev/null | grep -i -A3 "stall\|Duration\|Issue Slots\|L2 Cache Throughput\|red\|atom" | head -120 > gpurun_out/prof_wg_details.txt
rm -f gpurun_out/prof_wg.ncu-rep
python - <<'PY'
import csv
rows = list(csv.reader(open("gpurun_out/prof_wg_raw.csv")))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ["gpu__time_duration.sum", "smsp__average_warps_issue_stalled", "smsp__warp_issue_stalled", "lts__t_sectors_op_red", "lts__t_sectors_op_atom", "l1tex__t_set_accesses_pipe_lsu_mem_global_op_red", "smsp__inst_executed_op_global_red", "sm__inst_executed.sum", "smsp__cycles_active.avg", "launch__grid_size", "lts__t_requests_srcunit_tex_op_red", "smsp__inst_executed.sum"]
for r in data:
    for i, h in enumerate(hdr):
        if any(k in h for k in keys):
            try: v = float(r[i].replace(",", ""))
            except ValueError: v = r[i]
            if v not in (0, 0.0, "0"): print(f"{h} [{units[i]}] = {r[i]}")
PY
