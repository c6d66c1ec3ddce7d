"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python tools/summarize_launches.py launches.csv > profiles/rNN_launches_X.txt"""
import csv, sys, re, collections
rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
agg = collections.OrderedDict()
total = 0.0
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += ms
    total += ms
print("# per-kernel totals of %s (cold-cache, serialised launches: shares matter, not absolutes)" % sys.argv[1])
print("%-60s %8s %12s %8s" % ("kernel", "launches", "total ms", "share"))
for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-60s %8d %12.3f %7.1f%%" % (name[:60], n, ms, 100 * ms / total))
print("%-60s %8d %12.3f" % ("TOTAL", sum(a[0] for a in agg.values()), total))
