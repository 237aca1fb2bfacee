#!/usr/bin/env python
"""DRAM bytes per push of k_advance_p from an ncu metrics CSV of the benchmarked command
(dram__bytes_read.sum + dram__bytes_write.sum per launch; ion launch first, electron launch second in every
step: advance() walks the species list head first) -> profiles/<tag>_fullsize_traffic.json, which bench.py
reads for roofline.traffic.  usage: traffic_from_ncu.py CSV PARTICLES_PER_SPECIES OUT.json"""
import csv, json, sys
path, np_sp, out = sys.argv[1], float(sys.argv[2]), sys.argv[3]
rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
ix = {h: i for i, h in enumerate(rows[0])}
per = {}
for r in rows[1:]:
    v = float(r[ix["Metric Value"]].replace(",", "")) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12,
                                                          "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(r[ix["Metric Unit"]], 1.0)
    per.setdefault(int(r[ix["ID"]]), {})[r[ix["Metric Name"]]] = v
ids = sorted(per)
launches = [{"id": i, "species": "ion" if k % 2 == 0 else "electron",
             "dram_bytes": per[i]["dram__bytes_read.sum"] + per[i]["dram__bytes_write.sum"],
             "ms_under_ncu": per[i].get("gpu__time_duration.sum")} for k, i in enumerate(ids)]
res = {"source": path, "particles_per_species": np_sp, "launches": launches}
for sp in ("ion", "electron"):
    b = [l["dram_bytes"] for l in launches if l["species"] == sp]
    res[sp + "_B_per_push"] = sum(b) / len(b) / np_sp
res["B_per_push"] = (res["ion_B_per_push"] + res["electron_B_per_push"]) / 2
json.dump(res, open(out, "w"), indent=1)
print({k: round(v, 2) for k, v in res.items() if k.endswith("per_push")})
