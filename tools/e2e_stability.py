#!/usr/bin/env python3
"""Repeat bench.py's host-input batch arm: P pipelines on device 0, R runs of K steps, wall clock per run."""
import os, sys, time
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import sift3d_b200 as s3
from sift3d_b200.synth import workload
P = int(sys.argv[1]) if len(sys.argv) > 1 else 4
R = int(sys.argv[2]) if len(sys.argv) > 2 else 25
K = int(sys.argv[3]) if len(sys.argv) > 3 else 8
NI = 4
vol, units = workload("C2_mri_256")
dets = [s3.Detector(0) for _ in range(P)]
pools = [[s3.Image.from_array(vol, units) for _ in range(NI)] for _ in range(P)]
kps = [s3.KeypointStore() for _ in range(P)]
dss = [s3.DescriptorStore() for _ in range(P)]
def run_batch(nsteps):
    n = nsteps * P
    ims = [pools[v % P][(v // P) % NI] for v in range(n)]
    s3.detect_describe_batch(dets, ims, kps=[kps[v % P] for v in range(n)], dss=[dss[v % P] for v in range(n)])
run_batch(2)
out = []
for r in range(R):
    t0 = time.perf_counter()
    run_batch(K)
    out.append(1e3 * (time.perf_counter() - t0) / (K * P))
print("ms per volume:", " ".join(f"{v:.2f}" for v in out))
