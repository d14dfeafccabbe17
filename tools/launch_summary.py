#!/usr/bin/env python
"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list: count, mean and total duration."""
import collections
import csv
import re
import sys


def summarise(path):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr, agg = None, collections.OrderedDict()
    for r in rows:
        if "Kernel Name" in r:
            hdr = r
            continue
        if hdr is None or len(r) < len(hdr):
            continue
        d = dict(zip(hdr, r))
        if d.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(d["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(d["Metric Unit"], 1.0)
        name = re.sub(r"^void\s+", "", re.sub(r"\(.*", "", d["Kernel Name"]))
        agg.setdefault(name[:72], []).append(v)
    return agg


if __name__ == "__main__":
    for f in sys.argv[1:]:
        print(f"== {f}")
        for k, v in summarise(f).items():
            print(f"{k:72s} n={len(v):4d} mean={sum(v) / len(v):9.1f} us total={sum(v):10.1f} us")
