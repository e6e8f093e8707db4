"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) of bench.py: per-kernel totals of
the last step and the per-launch list of one chunk."""
import csv, collections, re, sys
path = sys.argv[1]
rows = [r for r in csv.reader(open(path)) if len(r) > 10]
hdr = rows[0]
ki, vi, gi = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Grid Size')
data = rows[1:]
idx = [i for i, r in enumerate(data) if 'minmax_f32' in r[ki]]
a = idx[-2] if len(idx) > 1 else 0
b = idx[-1]
step = data[a:b]
def short(n):
    m = re.search(r'(\w+_kernel)(<[^>]*>)?', n)
    return (m.group(1) + (m.group(2) or '')) if m else n[:40]
us = lambda r: float(r[vi].replace(',', '')) / 1e3
tot = sum(us(r) for r in step)
print(f'step total {tot:.1f} us, {len(step)} launches')
agg = collections.OrderedDict()
for r in step:
    k = short(r[ki]); agg.setdefault(k, [0, 0.0]); agg[k][0] += 1; agg[k][1] += us(r)
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f'{k:40s} n={c:4d} t={t:9.1f}us {100*t/tot:5.1f}%')
stems = [i for i, r in enumerate(step) if 'stem_kernel' in r[ki]]
if len(stems) > 1:
    ch = step[stems[0]:stems[1]]
    print(f'--- one chunk: {sum(us(r) for r in ch):.1f} us')
    for r in ch:
        print(f'{short(r[ki]):36s} grid={r[gi]:>14s} {us(r):8.1f}')
