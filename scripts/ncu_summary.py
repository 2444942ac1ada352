"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count / mean / total
and the launch sequence of one ba() call.  usage: python scripts/ncu_summary.py launches.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
cols = rows[hdr]
ki, vi = cols.index("Kernel Name"), cols.index("Metric Value")
agg, order = collections.defaultdict(list), []
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    name = r[ki].split("(")[0].replace("dbaf::<unnamed>::", "").replace("void ", "")[-58:]
    v = float(r[vi].replace(",", ""))
    agg[name].append(v)
    order.append((name, v))
print(f"{'kernel':60s} {'n':>4s} {'mean us':>9s} {'total us':>10s}")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k:60s} {len(v):4d} {sum(v) / len(v) / 1000:9.2f} {sum(v) / 1000:10.1f}")
idx = [i for i, (n, v) in enumerate(order) if "k_tables" in n]
if idx:
    i0 = idx[len(idx) // 2]
    print("--- one ba() call (2 iterations):")
    tot = 0
    for n, v in order[i0:]:
        if "k_tables" in n and tot > 0:
            break
        print(f"   {n:58s} {v / 1000:8.2f} us")
        tot += v
    print(f"   {'sum':58s} {tot / 1000:8.2f} us")
