#!/usr/bin/env python
"""Shared-memory wavefronts per SASS instruction of an .ncu-rep: where the L1 data pipe's time goes.
usage: tools/ncu_smem.py <rep> [n]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = next(k for k, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r)
hdr = rows[h]
isrc, iex = hdr.index('Source'), hdr.index('Instructions Executed')
iw, ii = hdr.index('L1 Wavefronts Shared'), hdr.index('L1 Wavefronts Shared Ideal')
L = []; byop = collections.Counter(); byop_i = collections.Counter(); byop_n = collections.Counter()
for k, r in enumerate(rows[h + 1:]):
    if len(r) <= iw or not r[iw].isdigit() or int(r[iw]) == 0: continue
    w, i, ex = int(r[iw]), int(r[ii] or 0), int(r[iex])
    op = r[isrc].split()[1] if r[isrc].startswith('@') else r[isrc].split()[0]
    L.append((w, i, ex, k, r[isrc])); byop[op] += w; byop_i[op] += i; byop_n[op] += ex
tot = sum(x[0] for x in L)
print('total shared wavefronts', tot, 'ideal', sum(x[1] for x in L))
for op, w in byop.most_common(): print(f'{op:14s} wavefronts {w:9d} ideal {byop_i[op]:9d} executed {byop_n[op]:9d}  wf/inst {w / byop_n[op]:.2f}')
for w, i, ex, k, src in sorted(L, reverse=True)[:n]: print(w, i, ex, k, src)
