"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals for the last step.

usage: python scripts/summarize_launches.py gpurun_out/launches.csv [marker_kernel_substring]
The step boundary is detected by the de-dup marker kernel (first kernel of every step)."""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    marker = sys.argv[2] if len(sys.argv) > 2 else "dedup_mark_kernel"
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((r["Kernel Name"], float(r["Metric Value"].replace(",", "")), r["Grid Size"], r["Block Size"]))
    starts = [i for i, r in enumerate(rows) if marker in r[0]]
    if len(starts) < 2:
        print("marker not found; summarising everything")
        seg = rows
    else:
        seg = rows[starts[-2]:starts[-1]]
    agg = OrderedDict()
    for name, ns, grid, block in seg:
        short = re.sub(r"\(.*", "", name)
        short = re.sub(r"^void ", "", short)[:90]
        a = agg.setdefault(short, [0, 0.0])
        a[0] += 1
        a[1] += ns
    total = sum(v[1] for v in agg.values())
    print(f"kernels in step: {len(seg)}  total device time: {total / 1e3:.1f} us")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{ns / 1e3:10.1f} us  {100 * ns / total:5.1f}%  x{n:<4d} {k}")


if __name__ == "__main__":
    main()
