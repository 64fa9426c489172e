#!/usr/bin/env python3
"""Executed warp-instructions and stall samples of one kernel of an .ncu-rep, per CUDA source
line (needs -lineinfo and --import-source on).
    python tools/ncu_lines.py rep.ncu-rep [kernel-name-regex] [top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
kid = sys.argv[2] if len(sys.argv) > 2 and sys.argv[2] else None
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]
if kid:
    cmd += ["--kernel-name", "regex:" + kid]
out = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, fpath, lines = None, "", []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        fpath = r[1].split("/")[-1]
    if len(r) > 8 and r[0] == "Line No":
        hdr = r
        ii, isamp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if hdr is None or len(r) <= ii or not r[0].isdigit():
        continue
    try:
        lines.append((fpath, int(r[0]), r[1].strip(), int(r[ii] or 0), int(r[isamp] or 0)))
    except ValueError:
        pass
ti, ts = sum(l[3] for l in lines), sum(l[4] for l in lines)
print(f"total warp-instructions {ti}, samples {ts}")
for f, ln, src, i, s in sorted(lines, key=lambda l: -l[3])[:top]:
    print(f"{100*i/max(1,ti):5.1f}% inst {100*s/max(1,ts):5.1f}% smp  {f}:{ln:<5d} {src[:95]}")
