#!/usr/bin/env python
"""Top SASS lines of an .ncu-rep by samples of one stall reason.  usage: tools/ncu_stalls.py <rep> <reason substring> [n]"""
import csv, io, subprocess, sys
rep, reason = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = next(k for k, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r)
hdr = rows[h]
cols = [i for i, c in enumerate(hdr) if c.startswith('stall_') and reason in c]
print([hdr[i] for i in cols])
isrc, iex = hdr.index('Source'), hdr.index('Instructions Executed')
L = []
for k, r in enumerate(rows[h + 1:]):
    if len(r) <= iex or not r[iex].isdigit(): continue
    v = sum(int(r[i]) for i in cols if r[i].isdigit())
    L.append((v, k, r[iex], r[isrc]))
tot = sum(v for v, *_ in L)
print('total', tot)
for v, k, ex, src in sorted(L, reverse=True)[:n]:
    print(v, k, ex, src)
