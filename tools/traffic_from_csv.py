"""ncu csv (gpu__time_duration.sum, dram__bytes_read.sum, dram__bytes_write.sum per launch) of ONE decode step ->
profiles JSON for bench.py's roofline.traffic + a per-kernel launch summary.
usage: python tools/traffic_from_csv.py traffic.csv step.json out.json"""
import collections
import csv
import json
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
H = rows[0]
ix = {h: i for i, h in enumerate(H)}
per = collections.defaultdict(lambda: collections.defaultdict(float))
ids = set()
for r in rows[1:]:
    if len(r) < len(H):
        continue
    name = r[ix["Kernel Name"]].split("(")[0].split("<")[0]
    val = float(r[ix["Metric Value"]].replace(",", "") or 0)
    unit = r[ix["Metric Unit"]]
    metric = r[ix["Metric Name"]]
    if metric == "gpu__time_duration.sum":
        val *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
    else:
        val *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
    per[name][metric] += val
    ids.add(r[ix["ID"]])
    per[name]["launches"] += 1.0 / 3
step = json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1])
g = [v for k, v in per.items() if "k_gemv" in k]
out = {"launches": len(ids), "gemv_launches": int(round(sum(v["launches"] for v in g))),
       "gemv_dram_bytes_read": sum(v["dram__bytes_read.sum"] for v in g), "gemv_dram_bytes_write": sum(v["dram__bytes_write.sum"] for v in g),
       "gemv_time_us_serialised": sum(v["gpu__time_duration.sum"] for v in g),
       "other_time_us_serialised": sum(v["gpu__time_duration.sum"] for k, v in per.items() if "k_gemv" not in k),
       "algorithmic_bytes_per_step": step["algorithmic_bytes_per_step"], "workload": step["workload"],
       "per_kernel": {k: {"launches": int(round(v["launches"])), "time_us": v["gpu__time_duration.sum"],
                          "dram_read": v["dram__bytes_read.sum"], "dram_write": v["dram__bytes_write.sum"]} for k, v in per.items()},
       "source": "ncu --nvtx --nvtx-include b200_step/ --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum "
                 "--clock-control none on B200 (cold-cache, serialised launches), one eager decode step of bench.py"}
json.dump(out, open(sys.argv[3], "w"), indent=1)
print(json.dumps({k: v for k, v in out.items() if k != "per_kernel"}))
