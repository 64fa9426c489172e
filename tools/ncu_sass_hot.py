#!/usr/bin/env python3
"""Per-instruction hot list of one kernel of an .ncu-rep (SASS view): executed count and
stall samples, grouped by opcode and the top individual instructions.

    python tools/ncu_sass_hot.py rep.ncu-rep 'regex:blur_fused:2' [top]
"""
import csv, io, subprocess, sys, collections
rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", "::" + kid],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1])
head = rows[1]
iS, iE, iW = head.index("Source"), head.index("Instructions Executed"), head.index("# Samples")
byop = collections.defaultdict(lambda: [0, 0, 0])
ins = []
for n, r in enumerate(rows[2:]):
    if len(r) <= iW: continue
    if r[iE] == 'Instructions Executed' or r[0] == 'Kernel Name': break
    src = r[iS].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    e, w = int(r[iE] or 0), int(r[iW] or 0)
    a = byop[op.split(".")[0]]
    a[0] += e; a[1] += w; a[2] += 1
    ins.append((w, e, n, src))
te = sum(a[0] for a in byop.values()); tw = sum(a[1] for a in byop.values())
print(f"total executed {te}  samples {tw}  static instrs {len(ins)}")
print("-- by opcode (executed share, stall-sample share)")
for op, a in sorted(byop.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"  {op:12s} exec {100*a[0]/te:5.1f}%  samples {100*a[1]/max(tw,1):5.1f}%  static {a[2]}")
print("-- top instructions by stall samples")
for w, e, n, src in sorted(ins, reverse=True)[:top]:
    print(f"  #{n:5d} samples {100*w/max(tw,1):5.2f}% exec {e:9d}  {src[:90]}")
