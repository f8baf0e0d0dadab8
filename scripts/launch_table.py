"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python scripts/launch_table.py list.csv"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
tot = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("ldc::", "").replace("void ", "")
    n, t = tot.get(name, (0, 0.0))
    tot[name] = (n + 1, t + float(r[14].replace(",", "")))
allt = sum(t for _, t in tot.values())
print(f"{len(rows)} launches, {allt / 1e6:.3f} ms of device time (cold, serialised under ncu)")
for name, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:40s} {n:6d} launches {t / 1e6:10.3f} ms {100 * t / allt:6.2f} %  {t / n / 1e3:10.2f} us each")
