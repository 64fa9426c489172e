#!/usr/bin/env python3
"""Host-side share of a step: wall time of detect / extract against their GPU stage totals."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200.synth import workload
vol, units = workload("C2_mri_256")
det = s3.Detector(0)
im = s3.Image.from_array(vol, units); im.make_resident(0)
kp, ds = s3.KeypointStore(), s3.DescriptorStore()
for _ in range(5):
    det.detect(im, kp); det.extract(kp, ds)
wd = we = gd = ge = 0.0
N = 50
for _ in range(N):
    t0 = time.perf_counter(); det.detect(im, kp); t1 = time.perf_counter(); det.extract(kp, ds); t2 = time.perf_counter()
    t = det.timings()
    wd += t1 - t0; we += t2 - t1; gd += t["total_detect"]; ge += t["total_describe"]
print(f"detect : wall {wd/N*1e3:.3f} ms, gpu {gd/N:.3f} ms, host-only {wd/N*1e3-gd/N:.3f} ms")
print(f"extract: wall {we/N*1e3:.3f} ms, gpu {ge/N:.3f} ms, host-only {we/N*1e3-ge/N:.3f} ms")
