"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel count / mean / share."""
import csv, sys, collections
lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
rows = list(csv.DictReader(lines))
d = collections.defaultdict(list)
for r in rows:
    d[r['Kernel Name'].split('(')[0].replace('void ', '')].append(float(r['Metric Value'].replace(',', '')))
tot = sum(sum(v) for v in d.values())
print("%-28s %5s %10s %9s %9s %6s" % ("kernel", "n", "mean_ns", "min_ns", "max_ns", "share"))
for k, v in d.items():
    print("%-28s %5d %10.0f %9.0f %9.0f %6.3f" % (k, len(v), sum(v) / len(v), min(v), max(v), sum(v) / tot))
