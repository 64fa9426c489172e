#!/usr/bin/env python3
"""Executed-instruction share by SASS region (runs of instructions with the same execution
count) of one kernel in an .ncu-rep:  python tools/ncu_regions.py rep.ncu-rep 'regex:name:1'"""
import csv, subprocess, sys, io
rep, kid = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", "::" + kid],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
head = rows[1]
iS, iE, iT = head.index('Source'), head.index('Instructions Executed'), head.index('Thread Instructions Executed')
iW = head.index('# Samples')
ins = []
for r in rows[2:]:
    if len(r) <= iT or r[iE] == 'Instructions Executed': continue
    ins.append((r[iS].strip(), int(r[iE] or 0), int(r[iT] or 0), int(r[iW] or 0)))
tot = sum(e for _, e, _, _ in ins); tw = sum(w for *_, w in ins)
print(f"total executed {tot}  samples {tw}")
i = 0
while i < len(ins):
    j = i
    while j + 1 < len(ins) and abs(ins[j + 1][1] - ins[i][1]) <= 0.02 * max(ins[i][1], 1): j += 1
    e = sum(x[1] for x in ins[i:j + 1]); t = sum(x[2] for x in ins[i:j + 1]); w = sum(x[3] for x in ins[i:j + 1])
    if e > 0.004 * tot:
        print(f'#{i:5d}-{j:5d} n={j-i+1:4d} exec/inst={ins[i][1]:9d} share={100*e/tot:5.1f}% samples={100*w/max(tw,1):5.1f}% lanes={t/max(e,1):5.1f}  {ins[i][0][:44]}')
    i = j + 1
