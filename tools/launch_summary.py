"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel launches, total / mean us, share."""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= vi: continue
    v = float(r[vi].replace(",", "")); u = r[ui]
    us = v / 1e3 if u in ("ns", "nsecond") else (v if u in ("us", "usecond") else v * 1e3 if u in ("ms", "msecond") else v)
    a = agg.setdefault(r[ki].split("(")[0][:60], [0, 0.0]); a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':60s} {'launches':>8s} {'total_us':>10s} {'mean_us':>9s} {'share':>6s}")
for k, (n, t) in agg.items():
    print(f"{k:60s} {n:8d} {t:10.1f} {t / n:9.2f} {t / tot:6.3f}")
