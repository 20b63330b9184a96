"""Aggregate an `ncu --csv` launch list by kernel and metric: python profiles/agg_launches.py file.csv"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = None; agg = collections.OrderedDict()
for r in rows:
    if 'Kernel Name' in r: hdr = r; continue
    if hdr is None or len(r) != len(hdr): continue
    d = dict(zip(hdr, r))
    k = (d['Kernel Name'][:34], d['Metric Name']); v = float(d['Metric Value'].replace(',', ''))
    a = agg.setdefault(k, [0, 0.0, d['Metric Unit']]); a[0] += 1; a[1] += v
for k, (n, t, u) in agg.items():
    print(f"{k[0]:36s} {k[1]:26s} n={n:4d} total={t:14.1f} {u} avg={t/n:12.1f}")
