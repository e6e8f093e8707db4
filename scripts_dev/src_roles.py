"""Per-region sample summary of one kernel in an `ncu --page source --csv --print-source sass` dump:
regions are split at the role markers (UTMALDG, UTCHMMA, LDTM, FFMA2) and the top stalled instructions are listed."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1]))); k = int(sys.argv[2]); minS = int(sys.argv[3]) if len(sys.argv) > 3 else 60
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name'] + [len(rows)]
a, b = starts[k], starts[k + 1]
hdr = rows[a + 1]; body = [r for r in rows[a + 2:b] if len(r) == len(hdr) and r[0].startswith('0x')]
S = hdr.index('# Samples'); SRC = hdr.index('Source'); IE = hdr.index('Instructions Executed')
sc = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[S]) for r in body)
print(rows[a][1][:60], 'samples', tot, 'instructions', len(body))
seg = 100
for s in range(0, len(body), seg):
    blk = body[s:s + seg]
    sm = sum(int(r[S]) for r in blk); ie = sum(int(r[IE] or 0) for r in blk)
    tags = sorted({key for r in blk for key in ('UTMALDG', 'UTCHMMA', 'LDTM', 'FFMA2', 'UTMASTG', 'UTMAREDG', 'VHMNMX', 'BAR.SYNC', 'SYNCS.PHASECHK', 'USETMAXREG') if key in r[SRC]})
    if sm > tot * 0.004: print(f'  [{s:5d}-{s+seg:5d}] samples {sm:6d} ({100*sm/tot:4.1f}%) inst_exec {ie/1e6:7.2f}M', tags)
for i, r in enumerate(body):
    if int(r[S]) >= minS:
        st = max(sc, key=lambda c: int(r[c] or 0))
        prev = body[i - 1][SRC].strip()[:58] if 'BRA' in r[SRC] else ''
        print(i, r[SRC].strip()[:64], r[S], r[IE], hdr[st], ('<- ' + prev) if prev else '')
