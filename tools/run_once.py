#!/usr/bin/env python3
"""One warm-up + one measured detect+describe on a named workload (for ncu captures)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200.synth import workload
name = sys.argv[1] if len(sys.argv) > 1 else "C2_mri_256"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
vol, units = workload(name)
det = s3.Detector(0)
im = s3.Image.from_array(vol, units)
kp, ds = s3.KeypointStore(), s3.DescriptorStore()
for _ in range(reps):
    det.detect(im, kp)
    det.extract(kp, ds)
print(json.dumps({"keypoints": len(kp), "timings": det.timings()}))
