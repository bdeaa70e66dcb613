"""Print the headline numbers of a bench.py JSON line (helper for gpurun one-liners)."""
import json
import sys

for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    r = d["roofline"]
    print(f, "ms/step", round(d["ms_per_step"], 4), "value", round(d["value"]), "gemm TF", round(r["achieved"], 2), "frac",
          round(r["frac"], 3), "share", round(r["share_of_step"], 3), "e2e", round(d["e2e"]["value"]))
