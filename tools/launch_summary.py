"""Summarise an ncu launch list (csv with gpu__time_duration.sum) per kernel."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hdr = None
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if "Kernel Name" in r:
        hdr = r; continue
    if hdr is None or len(r) != len(hdr):
        continue
    d = dict(zip(hdr, r))
    try:
        v = float(d["Metric Value"].replace(",", ""))
    except ValueError:
        continue
    if d.get("Metric Unit", "ns") in ("us", "usecond"):
        v *= 1e3
    n = d["Kernel Name"][:58]
    agg[n][0] += 1; agg[n][1] += v
tot = sum(t for _, t in agg.values())
print("%-58s %7s %12s %10s %6s" % ("kernel", "count", "total us", "avg us", "share"))
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-58s %7d %12.1f %10.2f %5.1f%%" % (n, c, t / 1e3, t / c / 1e3, 100 * t / tot))
