"""profiles/traffic.json from an `ncu --set full` capture of tools/stage_bench.py (one launch per
stage): measured DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per stage call.
    ncu -i rep.ncu-rep --page raw --csv > raw.csv ; python tools/traffic_from_ncu.py raw.csv [out.json]"""
import csv
import json
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stage_of = [("ssd_rows_kernel", "ssd"), ("ssd_tiled_kernel", "ssd"), ("birchfield_row", "bt"), ("sgm_", "sgm"),
            ("wta_vec_kernel", "wta"), ("dp_row", "dp")]
per_kernel, stage = [], {}
after_dp = False
seen = {}
for r in rows[2:]:
    name = r[ix["Kernel Name"]]
    rd = float(r[ix["dram__bytes_read.sum"]]) * UNIT[units[ix["dram__bytes_read.sum"]]]
    wr = float(r[ix["dram__bytes_write.sum"]]) * UNIT[units[ix["dram__bytes_write.sum"]]]
    t_us = float(r[ix["gpu__time_duration.sum"]]) * {"us": 1.0, "ms": 1e3, "ns": 1e-3}[units[ix["gpu__time_duration.sum"]]]
    st = next((s for k, s in stage_of if k in name), None)
    if st == "dp":
        after_dp = True
    elif st == "sgm" and after_dp:
        st = "sgm_wta"  # stage_bench's last stage: aggregation with the argmin folded in, no volume stored
    per_kernel.append({"kernel": name[:110], "full_name": name, "stage": st, "us": round(t_us, 1), "dram_read": rd, "dram_write": wr})
    if st is None:
        continue
    key = (st, name)
    seen[key] = seen.get(key, 0) + 1
# stage_bench launches ssd (cost), bt, ssd (restore), sgm, wta, dp, sgm_wta once each: a stage's traffic is the
# sum over the kernels of ONE call; kernels that were captured twice (ssd) are averaged
acc = {}
for k in per_kernel:
    if k["stage"] is None:
        continue
    n = seen[(k["stage"], k["full_name"])]
    acc[k["stage"]] = acc.get(k["stage"], 0.0) + (k["dram_read"] + k["dram_write"]) / n
out = {s: round(v) for s, v in acc.items()}
out["_source"] = "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per stage call at 1080x1920x128"
for k in per_kernel:
    del k["full_name"]
out["_kernels"] = per_kernel
path = sys.argv[2] if len(sys.argv) > 2 else None
text = json.dumps(out, indent=1)
if path:
    open(path, "w").write(text + "\n")
print(json.dumps({k: v for k, v in out.items() if not k.startswith("_")}))
