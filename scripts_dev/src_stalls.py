"""Aggregate an `ncu --page source --csv` dump: per kernel, stall-reason totals for a SASS address range
(selected by an opcode filter) and the top instructions by samples."""
import csv, sys, re, collections
rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
# split kernels
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
starts.append(len(rows))
k = 0
seen = 0
for a, b in zip(starts[:-1], starts[1:]):
    hdr = rows[a + 1]
    body = [r for r in rows[a + 2:b] if len(r) == len(hdr) and r[0].startswith('0x')]
    if not body: continue
    if seen != which: seen += 1; continue
    seen += 1
    print(rows[a][1][:60], 'instr rows', len(body))
    S = hdr.index('# Samples'); SRC = hdr.index('Source'); IE = hdr.index('Instructions Executed')
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    # classify rows by region: find the dw region = between first FFMA2/FMUL2 and last
    idx = [i for i, r in enumerate(body) if re.search(r'FFMA2|FMUL2', r[SRC])]
    lo, hi = (min(idx), max(idx)) if idx else (0, 0)
    tot = collections.Counter(); n = 0; ie = 0
    opc = collections.Counter(); opc_s = collections.Counter()
    for r in body[max(0, lo - 40):hi + 40]:
        for c in stall_cols: tot[hdr[c]] += int(r[c] or 0)
        n += int(r[S] or 0); ie += int(r[IE] or 0)
        m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[SRC])
        op = m.group(2).split('.')[0] if m else '?'
        opc[op] += int(r[IE] or 0); opc_s[op] += int(r[S] or 0)
    allS = sum(int(r[S] or 0) for r in body)
    print('dw region rows', lo, hi, 'samples', n, 'of', allS, 'inst', ie)
    for kx, v in tot.most_common(8): print('   %-28s %7d %5.1f%%' % (kx, v, 100.0 * v / max(1, sum(tot.values()))))
    print('   opcodes (inst executed, samples):')
    for kx, v in opc.most_common(16): print('     %-10s %10d %5.1f%%  samples %6d' % (kx, v, 100.0 * v / max(1, ie), opc_s[kx]))
