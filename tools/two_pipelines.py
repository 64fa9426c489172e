#!/usr/bin/env python3
"""Throughput of P detectors on ONE device, each on its own host thread (ctypes releases the GIL),
against one detector: do the kernels of independent pipelines fill each other's tails?
    python tools/two_pipelines.py [P] [volumes per pipeline]"""
import os, sys, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200.synth import workload
P = int(sys.argv[1]) if len(sys.argv) > 1 else 2
K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
vol, units = workload("C2_mri_256")
def run(p, k, out):
    det = s3.Detector(0)
    im = s3.Image.from_array(vol, units)
    kp, ds = s3.KeypointStore(), s3.DescriptorStore()
    dev = im.upload(0) if hasattr(im, "upload") else None
    for _ in range(3):
        det.detect(im, kp); det.extract(kp, ds)
    bar.wait()
    t0 = time.perf_counter()
    for _ in range(k):
        det.detect(im, kp); det.extract(kp, ds)
    out[p] = (time.perf_counter() - t0, len(kp))
for p in (1, P):
    bar = threading.Barrier(p)
    out = {}
    th = [threading.Thread(target=run, args=(i, K, out)) for i in range(p)]
    [t.start() for t in th]; [t.join() for t in th]
    wall = max(v[0] for v in out.values())
    print(f"{p} pipeline(s): {1e3 * wall / (K * p):.3f} ms per volume (host->device copy included), keypoints {out[0][1]}")
