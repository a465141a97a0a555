"""Per-source-line instruction counts of the first kernel in an ncu report.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass | python profiles/src_hot.py [N_TOP] [file-filter]"""
import csv, sys
rows = list(csv.reader(sys.stdin))
top = int(sys.argv[1]) if len(sys.argv) > 1 else 40
filt = sys.argv[2] if len(sys.argv) > 2 else ""
cur, hdr, agg, nfun = None, None, [], 0
for r in rows:
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r and r[0] == "Function Name":
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < 10 or r[2] != "-":
        continue
    try:
        ln = int(r[0]); n = int(r[hdr.index("Instructions Executed")]); s = int(r[hdr.index("# Samples")])
    except ValueError:
        continue
    agg.append((n, s, cur, ln, r[1].strip()[:100]))
tot = sum(a[0] for a in agg); ts = sum(a[1] for a in agg)
print("warp instructions", tot, "samples", ts)
for n, s, f, ln, src in sorted(agg, reverse=True)[:top]:
    if filt in f:
        print("%8d %5.1f%%  smp %5.1f%%  %s:%d  %s" % (n, 100 * n / tot, 100 * s / max(ts, 1), f, ln, src))
