"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name: python tools/ncu_agg.py file.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
agg = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(',', ''))
    v = v / 1000 if r[ui] == 'ns' else v * 1000 if r[ui] == 'ms' else v
    agg.setdefault(r[ki][:70], []).append(v)
tot = sum(sum(v) for v in agg.values())
for k, v in agg.items():
    print(f'{k:70s} n={len(v):5d} sum={sum(v):10.1f}us ({100 * sum(v) / tot:5.1f}%) avg={sum(v) / len(v):8.1f} max={max(v):8.1f}')
print('total us', round(tot, 1))
