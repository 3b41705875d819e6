#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: one training step (from after the previous
optimizer kernel to the optimizer kernel), grouped by kernel (GEMMs by template instantiation)."""
import collections
import csv
import re
import sys


def main(path, md=False):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        us = v * {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
        rows.append((row["Kernel Name"], us))
    ends = [i for i, x in enumerate(rows) if "adam_flat" in x[0] or "sharded_adam" in x[0]]
    if len(ends) >= 2:
        step = rows[ends[-2] + 1:ends[-1] + 1]
    else:
        step = rows
    tot = sum(x[1] for x in step)
    agg = collections.defaultdict(lambda: [0, 0.0])
    for k, t in step:
        k = re.sub(r"\(.*", "", k).replace("b200::", "").replace("void ", "")
        if "gemm_tcgen05" in k:
            m = re.search(r"<(.*)>", k)
            k = "gemm<" + m.group(1).replace("__nv_bfloat16", "bf16").replace("(bool)", "").replace(" ", "") + ">"
        elif k.startswith("at::"):
            k = "torch elementwise / copy kernels"
        agg[k][0] += 1
        agg[k][1] += t
    print("%d launches captured, one step = %d launches, %.1f us summed (serialised, cold-cache: compare shares)" % (len(rows), len(step), tot))
    if md:
        print("\n| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if md:
            print("| `%s` | %d | %.1f | %.1f%% | %.1f |" % (k, n, t, 100 * t / tot, t / n))
        else:
            print("%-64s %4d %9.1f us %5.1f%%  avg %7.1f" % (k[:64], n, t, 100 * t / tot, t / n))


if __name__ == "__main__":
    main(sys.argv[1], md="--md" in sys.argv)
