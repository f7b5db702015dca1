#!/usr/bin/env python
"""profiles/traffic.json from the ncu CSV written by scripts/gpu_traffic.sh:
python scripts/make_traffic_json.py gpurun_out/<tag>/traffic_512.csv"""
import csv, json, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
h = rows[0]; ci = {n: i for i, n in enumerate(h)}
per = {}
for r in rows[1:]:
    if r[0] == "ID": continue
    k = r[ci["ID"]]; per.setdefault(k, {})[r[ci["Metric Name"]]] = (float(r[ci["Metric Value"]].replace(",", "")), r[ci["Metric Unit"]])
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
launches = []
for k, m in per.items():
    b = sum(m[n][0] * scale[m[n][1]] for n in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    launches.append(b)
out = {"source": sys.argv[1], "kernel": "rhs_update_v5 (substeps 1,2,3 of one 512^3 step)", "dram_bytes_per_launch": launches,
       "dram_bytes_per_launch_512": sum(launches) / len(launches)}
json.dump(out, open("profiles/traffic.json", "w"), indent=1)
print(out)
