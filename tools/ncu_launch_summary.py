"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import sys, csv, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith('=='))]
H = rows[0]
ik, iv = H.index('Kernel Name'), H.index('Metric Value')
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= iv: continue
    n = r[ik].split('(')[0].replace('orbfe::<unnamed>::', '').replace('void ', '')
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += float(r[iv].replace(',', ''))
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':50s} {'launches':>8s} {'total us':>12s} {'share':>7s}")
for n, a in agg.items():
    print(f"{n[:50]:50s} {a[0]:8d} {a[1] / 1e3:12.1f} {a[1] / tot * 100:6.1f}%")
