#!/usr/bin/env python3
"""Upload time (stage 'h2d') of the 128^3 CLI configuration, rep by rep."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200.synth import WORKLOADS, workload
name = "C1_blobs_128"
(nx, ny, nz), units, _, _ = WORKLOADS[name]
for trial in range(3):
    im = s3.Image(nx, ny, nz)
    workload(name, out=im.array)
    im.set_units(*units)
    det = s3.Detector(0)
    kp, ds = s3.KeypointStore(), s3.DescriptorStore()
    h, w = [], []
    for i in range(12):
        t0 = time.perf_counter()
        det.detect(im, kp)
        kp.sort_by_strength(100)
        det.extract(kp, ds)
        w.append(1e3 * (time.perf_counter() - t0))
        h.append(det.timings()["h2d"])
    print("trial", trial, "h2d ms:", " ".join(f"{v:.2f}" for v in h), "| wall:", " ".join(f"{v:.2f}" for v in w))
    det.close()
