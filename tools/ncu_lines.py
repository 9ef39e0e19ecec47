#!/usr/bin/env python
"""Where a kernel's instructions go: SASS lines of an .ncu-rep (source page) grouped by execution count.
usage: python tools/ncu_lines.py <report.ncu-rep> [min share %] ; prints groups, then (with -v) every line."""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = next(k for k, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r)
hdr = rows[h]
isrc, iex, ismp = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
lines = [(r[isrc], int(r[iex]), int(r[ismp] or 0)) for r in rows[h + 1:] if len(r) > iex and r[iex].isdigit()]
tot = sum(n for _, n, _ in lines); ts = sum(s for _, _, s in lines) or 1
# consecutive runs with (nearly) the same execution count
runs = []
for i, (src, n, s) in enumerate(lines):
    if runs and abs(runs[-1][2] - n) <= max(2, 0.02 * n):
        runs[-1][1] = i; runs[-1][3] += n; runs[-1][4] += s
    else:
        runs.append([i, i, n, n, s])
print(f"total {tot} warp instructions, {len(lines)} SASS lines")
for a, b, n, sumn, s in runs:
    if sumn / tot >= float(sys.argv[2]) / 100 if len(sys.argv) > 2 and sys.argv[2] != '-v' else sumn / tot >= 0.005:
        ops = collections.Counter((l[0].split()[1] if l[0].startswith('@') else l[0].split()[0]).split('.')[0] for l in lines[a:b + 1])
        print(f"lines {a:5d}-{b:5d} ({b - a + 1:4d}) x {n:7d} = {100 * sumn / tot:5.1f}% instr {100 * s / ts:5.1f}% samples  " + " ".join(f"{o}:{c}" for o, c in ops.most_common(8)))
if '-v' in sys.argv:
    for i, (src, n, s) in enumerate(lines):
        print(i, n, s, src)
