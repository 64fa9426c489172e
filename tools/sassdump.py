#!/usr/bin/env python3
"""Instructions of one kernel between two SASS addresses, optionally without some opcodes.
    python tools/sassdump.py obj mangled-name lo hi [OP1,OP2,...]"""
import re, subprocess, sys
obj, fun, lo, hi = sys.argv[1], sys.argv[2], int(sys.argv[3], 16), int(sys.argv[4], 16)
skip = sys.argv[5].split(',') if len(sys.argv) > 5 else []
out = subprocess.run(["cuobjdump", "-sass", "-fun", fun, obj], capture_output=True, text=True).stdout
for line in out.splitlines():
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if not m:
        continue
    a = int(m.group(1), 16)
    if a < lo or a > hi:
        continue
    t = m.group(2).strip()
    op = re.sub(r"^@!?U?P\d\s+", "", t).split()[0].split('.')[0]
    if op in skip:
        continue
    print(f"{a:05x} {t}")
