"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, mean, share."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
H = rows[hdr]
ki, vi = H.index('Kernel Name'), H.index('Metric Value')
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
agg = collections.defaultdict(list)
for r in rows[hdr + 1 + skip:]:
    if len(r) > vi:
        try:
            agg[r[ki][:64]].append(float(r[vi].replace(',', '')))
        except ValueError:
            pass
tot = sum(sum(v) for v in agg.values())
print(f"{'kernel':66s} {'n':>5s} {'avg us':>9s} {'share':>7s}")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k:66s} {len(v):5d} {sum(v) / len(v) / 1e3:9.2f} {sum(v) / tot * 100:6.1f}%")
