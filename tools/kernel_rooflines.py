"""Per-kernel roofline table from `ncu -i X.ncu-rep --page raw --csv` exports (one or more files).

    python tools/kernel_rooflines.py label1=raw1.csv [label2=raw2.csv ...] > profiles/r02_kernel_rooflines.md

For every launch in the exports: duration, DRAM bytes moved and the fraction of the HBM peak that is
(MEASURED_PEAKS.json `hbm_gbps` when present, else 6550 GB/s), the FP64 pipe (DFMA) and DMMA sub-pipe activity, issue
activity, resident warps, and which of those bounds the launch ("latency" when no unit is above 40 %).
Fractions: DRAM = bytes / duration / peak; FP64 / DMMA = ncu's pct_of_peak_sustained_active for the pipe.
"""
import csv
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def hbm_peak_gbps() -> float:
    try:
        with open(os.path.join(HERE, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        for k in ("hbm_gbs", "hbm_gbps", "hbm_copy_gbps"):
            if k in d:
                return float(d[k])
        for v in d.values():
            if isinstance(v, dict):
                for k, x in v.items():
                    if "hbm" in k.lower() and isinstance(x, (int, float)):
                        return float(x)
    except Exception:
        pass
    return 6550.0


UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def short(name: str) -> str:
    name = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", name)
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*\)$", "", name)
    return name[:88]


def main() -> None:
    peak = hbm_peak_gbps()
    print(f"HBM peak used: {peak:.0f} GB/s.  One row per captured launch (`ncu --set full`, cold cache, serialised).\n")
    for arg in sys.argv[1:]:
        label, path = arg.split("=", 1)
        rows = list(csv.reader(open(path, errors="replace")))
        hdr, units = rows[0], rows[1]
        col = {h: i for i, h in enumerate(hdr)}

        def get(r, name, default=0.0):
            i = col.get(name)
            if i is None or r[i] in ("", "n/a"):
                return default
            try:
                return float(r[i].replace(",", "")) * UNIT.get(units[i], 1.0)
            except ValueError:
                return default

        print(f"### {label}\n")
        print("| kernel | grid x block | regs | us | DRAM MB | DRAM frac | FP64 pipe % | DMMA % | issue % | warps % | bound |")
        print("|---|---|---|---|---|---|---|---|---|---|---|")
        for r in rows[2:]:
            if len(r) < len(hdr) // 2:
                continue
            us = get(r, "gpu__time_duration.sum")
            if us <= 0:
                continue
            byts = get(r, "dram__bytes_read.sum") + get(r, "dram__bytes_write.sum")
            dfrac = byts / (us * 1e-6) / (peak * 1e9)
            fp64 = get(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active")
            dmma = get(r, "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active")
            issue = get(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")
            warps = get(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
            l2 = get(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed")
            cands = {"hbm": 100 * dfrac, "fp64": fp64, "dmma": dmma, "issue": issue, "l2": l2}
            top = max(cands, key=cands.get)
            bound = f"{top} {cands[top]:.0f}%" if cands[top] >= 40 else f"latency (top: {top} {cands[top]:.0f}%)"
            grid = f"{int(get(r, 'launch__grid_size'))} x {int(get(r, 'launch__block_size'))}"
            print(f"| `{short(r[col['Kernel Name']])}` | {grid} | {int(get(r, 'launch__registers_per_thread'))} | {us:.1f} | "
                  f"{byts / 1e6:.1f} | {dfrac:.2f} | {fp64:.0f} | {dmma:.0f} | {issue:.0f} | {warps:.0f} | {bound} |")  # fmt: skip
        print()


if __name__ == "__main__":
    main()
