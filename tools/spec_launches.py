"""Per-kernel times of the last samples-variant pass in an ncu launch list (gpu__time_duration.sum CSV)."""
import csv, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith('=='))]
h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value'); ui = h.index('Metric Unit')
rs = rows[1:]
last = max(i for i, r in enumerate(rs) if 'Primary' in r[ki])
tot = 0
for r in rs[last:]:
    v = float(r[vi].replace(',', '')); u = r[ui]
    ms = v / 1e6 if u in ('ns', 'nsecond') else v / 1e3 if u in ('us', 'usecond') else v
    tot += ms
    print('  %-40s %.3f ms' % (r[ki][:40], ms))
print('  total %.3f' % tot)
