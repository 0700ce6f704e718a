"""Summarise `ncu -i X.ncu-rep --page source --csv --kernel-name ... --launch-count 1`: executed
instructions per opcode, and contiguous SASS regions with equal execution count (loop bodies)."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r)
hdr = rows[h]
iS, iE, iW, iP = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('L1 Wavefronts Shared'), hdr.index('# Samples')
tot, byop, wf, samp, seq = 0, collections.Counter(), collections.Counter(), collections.Counter(), []
for r in rows[h + 1:]:
    if len(r) <= iE or not r[iE].isdigit():
        continue
    src, n = r[iS].strip(), int(r[iE])
    op = re.sub(r'^@!?U?P\d+\s+', '', src).split()[0].split('.')[0]
    s = int(r[iP] or 0)
    tot += n
    byop[op] += n
    wf[op] += int(r[iW] or 0)
    samp[op] += s
    seq.append((n, src, s))
print('total warp instructions', tot)
for op, n in byop.most_common(24):
    print(f"{op:8s} {n:>11d} {100 * n / tot:5.1f}%  smem_wavefronts={wf[op]:>10d} stall_samples={samp[op]}")
reg, cur = [], None
for n, src, s in seq:
    if cur and cur[0] == n:
        cur[1] += 1
        cur[2] += s
    else:
        cur = [n, 1, s, src]
        reg.append(cur)
print('regions (exec count, #instrs, share of executed, stall samples, first instr)')
for r in reg:
    if r[0] * r[1] > 0.004 * tot:
        print(f"{r[0]:>9d} {r[1]:>5d} {100 * r[0] * r[1] / tot:5.1f}% {r[2]:>6d}  {r[3][:70]}")
