"""Top sampled SASS instructions of kernel #k in an `ncu --page source --csv` dump, with the wait target."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]))); k = int(sys.argv[2]); n = int(sys.argv[3]) if len(sys.argv) > 3 else 20
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name'] + [len(rows)]
a, b = starts[k], starts[k + 1]
hdr = rows[a + 1]; body = [r for r in rows[a + 2:b] if len(r) == len(hdr) and r[0].startswith('0x')]
S = hdr.index('# Samples'); SRC = hdr.index('Source'); IE = hdr.index('Instructions Executed')
sc = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
print(rows[a][1][:70], 'total samples', sum(int(r[S]) for r in body))
top = sorted(range(len(body)), key=lambda i: -int(body[i][S] or 0))[:n]
for i in sorted(top):
    r = body[i]; st = max(sc, key=lambda c: int(r[c] or 0))
    extra = ('   <- ' + body[i - 1][SRC].strip()[:60]) if 'BRA' in r[SRC] else ''
    print(i, r[SRC].strip()[:60], r[S], r[IE], hdr[st], extra)
