#!/usr/bin/env python
"""tools/ncu_traffic.py <report.ncu-rep> <kernel-substring> <json-key> <field> <units> — take dram__bytes_read.sum +
dram__bytes_write.sum of the longest launch of the named kernel in an `ncu --set full` report (read on the CPU box with `ncu -i`)
and store it PER UNIT of work (`units` = queries of that launch, or queries x rows for the scan) in profiles/ncu_traffic.json,
which bench.py reads for roofline.traffic. Example:
    python tools/ncu_traffic.py gpurun_out/r2_hnsw.ncu-rep hnsw_search_kernel hnsw_search_kernel dram_bytes_per_query 592"""
import csv, json, os, subprocess, sys
rep, sub, key, field, units = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4], float(sys.argv[5])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, unit_row = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
best = None
for r in rows[2:]:
    if sub in r[col["Kernel Name"]]:
        t = float(r[col["gpu__time_duration.sum"]].replace(",", ""))
        if best is None or t > best[0]:
            best = (t, r)
assert best, "no launch of %s in %s" % (sub, rep)
r = best[1]
tot = 0.0
for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
    tot += float(r[col[m]].replace(",", "")) * SCALE[unit_row[col[m]]]
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
try:
    db = json.load(open(path))
except Exception:
    db = {}
db[key] = {field: tot / units, "dram_bytes_per_launch": tot, "units_per_launch": units,
           "kernel": r[col["Kernel Name"]][:120], "launch_duration": r[col["gpu__time_duration.sum"]] + " " + unit_row[col["gpu__time_duration.sum"]],
           "source": "profiles/%s (ncu --set full --clock-control none of the shipped kernel)" % os.path.basename(rep).replace(".ncu-rep", "_ncu_full.txt")}
json.dump(db, open(path, "w"), indent=1, sort_keys=True)
print(json.dumps(db[key]))
