#!/usr/bin/env python3
"""Stall-reason totals of one kernel in an .ncu-rep and the instructions that own them.
    python tools/ncu_stalls.py rep.ncu-rep 'regex:name:N' [top]"""
import csv, io, subprocess, sys, collections
rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 12
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", "::" + kid],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1])
head = rows[1]
iS = head.index("Source")
cols = [i for i, h in enumerate(head) if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter(); per = []
for n, r in enumerate(rows[2:]):
    if len(r) <= max(cols) or r[0] == "Address" or r[0] == "Kernel Name": break
    d = {head[i]: int(r[i] or 0) for i in cols}
    tot.update(d); per.append((n, r[iS].strip(), d))
T = sum(tot.values())
print("total samples", T)
for k, v in tot.most_common():
    if v: print(f"  {k:24s} {100*v/T:5.1f}%")
for reason in [k for k, v in tot.most_common(4)]:
    print(f"-- top owners of {reason}")
    for n, src, d in sorted(per, key=lambda x: -x[2][reason])[:top]:
        if d[reason]: print(f"   #{n:5d} {100*d[reason]/T:5.2f}%  {src[:80]}")
