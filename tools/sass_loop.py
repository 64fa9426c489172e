#!/usr/bin/env python3
"""Static view of a kernel's SASS loops: for every backward branch, the instruction count of
the loop body and an opcode histogram.  python tools/sass_loop.py lib.so 'mangled-substring'"""
import collections, re, subprocess, sys
lib, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
blocks = out.split("Function : ")
for blk in blocks[1:]:
    name = blk.split("\n", 1)[0]
    if pat not in name: continue
    ins = []
    for line in blk.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
    addr2i = {a: i for i, (a, _) in enumerate(ins)}
    print(name, len(ins), "instructions")
    for i, (a, t) in enumerate(ins):
        m = re.search(r"BRA(?:\.U)?(?:\.ANY)?\s+(?:!?U?P\d,\s+)?(0x[0-9a-f]+)", t)
        if not m: continue
        tgt = int(m.group(1), 16)
        if tgt < a and tgt in addr2i:
            body = ins[addr2i[tgt]: i + 1]
            if len(body) < 40: continue
            h = collections.Counter()
            for _, s in body:
                s = re.sub(r"^@!?U?P\d\s+", "", s)
                h[s.split()[0].split(".")[0]] += 1
            print(f"  loop {tgt:#x}..{a:#x}: {len(body)} instrs: " +
                  " ".join(f"{k}:{v}" for k, v in h.most_common(24)))
