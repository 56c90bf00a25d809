#!/usr/bin/env python
"""DRAM traffic per launch of one kernel from the raw page of an `ncu --set full` capture:
    python tools/traffic_from_ncu.py gpurun_out/prof_step_raw.csv conv_tc_kernel profiles/conv_tc_traffic.json "<capture command>"
bench.py copies `bytes_per_launch` into roofline.traffic (dram__bytes_read.sum + dram__bytes_write.sum, averaged over the
captured launches of the kernel)."""
import csv, json, sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
path, kernel, out, cmd = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else ""
rows = list(csv.reader(open(path)))
hdr, units, data = rows[0], rows[1], rows[2:]
kn, rd, wr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
tot, n = 0.0, 0
for r in data:
    if kernel in r[kn]:
        tot += float(r[rd].replace(",", "")) * UNIT[units[rd]] + float(r[wr].replace(",", "")) * UNIT[units[wr]]
        n += 1
json.dump({"kernel": kernel, "launches": n, "bytes_per_launch": tot / max(n, 1), "source": path, "capture": cmd,
           "metric": "dram__bytes_read.sum + dram__bytes_write.sum"}, open(out, "w"), indent=1)
print(out, n, tot / max(n, 1))
