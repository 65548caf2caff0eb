"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): time and launches per kernel name,
and every kernel's share of the summed GPU time.  usage: python tools/sumlaunch.py profiles/rNN_launches_bench.csv"""
import csv, re, sys
from collections import defaultdict

def main(path):
    rows = [l for l in open(path, newline="") if l.startswith('"')]
    rd = csv.DictReader(rows)
    t, n = defaultdict(float), defaultdict(int)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum": continue
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1e-3)
        t[name] += v; n[name] += 1
    tot = sum(t.values())
    print("%-60s %8s %12s %8s %7s" % ("kernel", "launches", "total us", "avg us", "share"))
    for k in sorted(t, key=t.get, reverse=True):
        print("%-60s %8d %12.1f %8.2f %6.1f%%" % (k[:60], n[k], t[k], t[k] / n[k], 100 * t[k] / tot))
    print("%-60s %8d %12.1f" % ("total", sum(n.values()), tot))

if __name__ == "__main__":
    try: main(sys.argv[1])
    except BrokenPipeError: pass
