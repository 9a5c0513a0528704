#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel launches, total ms, share."""
import csv, sys, collections, re
def main(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(lines)
    tot = collections.OrderedDict()
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
        c = tot.setdefault(name, [0, 0.0]); c[0] += 1; c[1] += ms
    total = sum(v[1] for v in tot.values())
    print("kernel,launches,total_ms,share")
    for k, (n, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print(f"{k},{n},{ms:.3f},{ms/total:.4f}")
if __name__ == "__main__":
    main(sys.argv[1])
