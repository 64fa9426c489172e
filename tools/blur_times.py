#!/usr/bin/env python3
"""Per-level blur times (CUDA events) of one detect on a named workload, all octaves.
SIFT3D_CUDA_SERIAL=1 gives clean per-kernel numbers (no overlap between octaves)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200.synth import workload
name = sys.argv[1] if len(sys.argv) > 1 else "C2_mri_256"
vol, units = workload(name)
det = s3.Detector(0)
det.set_profiling(True)
im = s3.Image.from_array(vol, units)
kp = s3.KeypointStore()
for _ in range(3):
    det.detect(im, kp)
o = 0
while True:
    try:
        row = []
        for s in range(-1 if o == 0 else 0, 5):
            ms, n = det.blur_ms(o, s)
            row.append(f"s={s}: {ms*1e3:7.1f}us/{n}")
        print(f"octave {o}: " + "  ".join(row))
        o += 1
    except Exception:
        break
print(json.dumps(det.timings()))
