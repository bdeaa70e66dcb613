"""Reduce an `ncu --page source --csv` dump to its hottest SASS lines (run on the GPU box: the full dump is too large to
bring back).  Usage: python tools/ncu_top_lines.py src.csv out.txt [n]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = None
for i, r in enumerate(rows):
    if any("Sampling" in c for c in r):
        hdr, body = r, rows[i + 1 :]
        break
if hdr is None:
    open(sys.argv[2], "w").write("no header found\n" + "\n".join(",".join(r) for r in rows[:5]))
    sys.exit(0)
n = int(sys.argv[3]) if len(sys.argv) > 3 else 60
samp = next(i for i, c in enumerate(hdr) if "Sampling" in c and "All" in c) if any("Sampling" in c and "All" in c for c in hdr) else next(i for i, c in enumerate(hdr) if "Sampling" in c)
src = next((i for i, c in enumerate(hdr) if c.strip() in ("Source", "SASS")), 1)


def val(r):
    try:
        return float(r[samp])
    except Exception:
        return 0.0


body = [r for r in body if len(r) == len(hdr)]
total = sum(val(r) for r in body) or 1.0
top = sorted(body, key=val, reverse=True)[:n]
with open(sys.argv[2], "w") as fh:
    fh.write("columns: " + " | ".join(hdr) + "\n")
    fh.write(f"total samples {total}\n")
    for r in top:
        fh.write(f"{100 * val(r) / total:6.2f}%  " + " | ".join(c[:60] for c in r) + "\n")
    # also the lines in address order around the main loop: every line with >= 0.3 % of the samples
    fh.write("\n-- in program order (>= 0.3 %) --\n")
    for r in body:
        if val(r) / total >= 0.003:
            fh.write(f"{100 * val(r) / total:6.2f}%  {r[0][:12]:12s} {r[src][:90]}\n")
