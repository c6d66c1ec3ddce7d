"""Digest of an `ncu --page source --csv` export (SASS view): opcode mix, segments by execution count, hottest stalls.
usage: python tools/sass_profile.py source.csv UNITS [kernel-index]   (UNITS = work units to normalise by, e.g. warp groups)"""
import csv, collections, re, sys
raw = list(csv.reader(open(sys.argv[1])))
G = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
# the export holds one table per captured launch: "Kernel Name" row, header row, instruction rows
tables, cur = [], None
for r in raw:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        tables.append(cur)
    elif cur is not None and len(r) > 5 and r[0].startswith("0x"):
        cur["rows"].append(r)
t = tables[which]
rows = t["rows"]
print("#", t["name"][:100], "(%d launches in the file)" % len(tables))
tot_s = sum(int(r[2] or 0) for r in rows) or 1
tot_i = sum(int(r[5] or 0) for r in rows)
print("# warp instructions per unit: %.1f   stall samples: %d" % (tot_i / G, tot_s))
byop, byi = collections.Counter(), collections.Counter()
for r in rows:
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[1])
    op = m.group(2).split(".")[0] if m else "?"
    byop[op] += int(r[2] or 0)
    byi[op] += int(r[5] or 0)
for op, s in byop.most_common(14):
    print("%-8s samples %5.1f%%  instr/unit %7.1f" % (op, 100 * s / tot_s, byi[op] / G))
print("# hottest instructions (index, share of samples, executions per unit)")
top = sorted(range(len(rows)), key=lambda k: -int(rows[k][2] or 0))[:30]
for k in sorted(top):
    r = rows[k]
    print("%5d %5.2f%% x%.2f  %s" % (k, 100 * int(r[2] or 0) / tot_s, int(r[5]) / G, r[1].strip()[:96]))
