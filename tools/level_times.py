#!/usr/bin/env python3
"""Per-level blur times (CUDA events, profiling pass) of a named workload: tools/level_times.py C3_ct_512x512x300"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200.synth import workload
name = sys.argv[1] if len(sys.argv) > 1 else "C3_ct_512x512x300"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
vol, units = workload(name)
det = s3.Detector(0)
im = s3.Image.from_array(vol, units)
kp = s3.KeypointStore()
for _ in range(2):
    det.detect(im, kp)
det.set_profiling(True)
acc = {}
for _ in range(reps):
    det.detect(im, kp)
    for o in range(8):
        for s in range(-1, 5):
            try:
                ms, n = det.blur_ms(o, s)
            except Exception:
                continue
            if n > 0:
                acc.setdefault((o, s), []).append((ms, n))
for (o, s), v in sorted(acc.items()):
    print(f"octave {o} level {s:2d}: {sum(m for m, _ in v) / len(v) * 1e3:8.1f} us  launches {v[0][1]}")
print(json.dumps({"timings": det.timings()}))
