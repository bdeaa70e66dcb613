"""Turn an `ncu --set full ... --page raw --csv` dump of one bench.py step into profiles/r02_traffic.json: DRAM bytes per
launch of the dominant kernel, tagged with the digest of the library sources it was captured from (bench.py reports the
figure only while the loaded library carries the same digest).  Usage: python tools/make_traffic_json.py raw.csv C3 512"""
import csv
import json
import os
import sys

sys.path.insert(0, ".")
from botorch_b200 import _lib  # noqa: E402

raw, cfg, B = sys.argv[1], sys.argv[2], int(sys.argv[3])
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
ki, ri, wi, ti = (hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
launches = []
for r in rows[2:]:
    if "dgemm_dmma_tma_kernel" in r[ki]:
        launches.append({"kernel": r[ki][:80], "dram_read_bytes": float(r[ri]) * scale[units[ri]],
                         "dram_write_bytes": float(r[wi]) * scale[units[wi]], "duration_us": float(r[ti]) * (1e-3 if units[ti] == "ns" else 1.0)})
assert launches, "no dominant-kernel launch in the capture"
mean = sum(l["dram_read_bytes"] + l["dram_write_bytes"] for l in launches) / len(launches)
path = os.path.join("profiles", "r02_traffic.json")
rec = json.load(open(path)) if os.path.exists(path) else {"workloads": {}}
rec["source_digest"] = _lib.lib().bo_source_digest().decode()
rec["capture"] = "profiles/r02_ncu_full_step.csv"
rec.setdefault("workloads", {})[f"{cfg}:{B}"] = {
    "dram_bytes_per_launch": mean, "launches": launches,
    "note": f"mean of dram__bytes_read.sum + dram__bytes_write.sum over the {len(launches)} dgemm_dmma_tma_kernel launches of one step; "
            "algorithmic operand bytes: forward A 36 + B 66 + C 74 MB, backward A 36 + B 74 + E 66 MB + partials 25 MB"}
json.dump(rec, open(path, "w"), indent=1)
print(json.dumps(rec["workloads"][f"{cfg}:{B}"])[:400])
