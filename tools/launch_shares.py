"""Kernel shares of an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file ...): per kernel name
the launch count, total and mean device time, share of the sum.  argv: csv [skip-first-N-launches]"""
import csv, sys, re, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
acc = collections.OrderedDict()
for r in rows[skip:]:
    name = re.sub(r"\(.*", "", r[4]).replace("h3d::", "").replace("void ", "")
    ns = float(r[14])
    a = acc.setdefault(name, [0, 0.0, r[7], r[8]])
    a[0] += 1; a[1] += ns
tot = sum(a[1] for a in acc.values())
print(f"# {sys.argv[1]}: {len(rows) - skip} launches, {tot / 1e6:.3f} ms of kernel time (cold-cache, serialised: compare SHARES, not absolutes)")
print(f"{'kernel':60s} {'launches':>8s} {'total ms':>10s} {'mean us':>9s} {'share':>7s}  block  grid")
for name, (n, t, blk, grd) in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    print(f"{name[:60]:60s} {n:8d} {t / 1e6:10.3f} {t / n / 1e3:9.2f} {100 * t / tot:6.1f}%  {blk} {grd}")
