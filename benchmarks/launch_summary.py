"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list:  python benchmarks/launch_summary.py FILE.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
t, n = collections.Counter(), collections.Counter()
for r in rows[1:]:
    v = float(r[vi].replace(',', ''))
    v *= {'ns': 1e-6, 'us': 1e-3, 'usecond': 1e-3, 'nsecond': 1e-6, 'ms': 1, 'msecond': 1}.get(r[ui], 1e-6)
    k = r[ki].split('(')[0][:50]
    t[k] += v
    n[k] += 1
tot = sum(t.values())
print(sys.argv[1], 'total %.1f ms' % tot)
for k, v in t.most_common(10):
    print('  %-50s %8.2f ms %5.1f%%  n=%d' % (k, v, 100 * v / tot, n[k]))
