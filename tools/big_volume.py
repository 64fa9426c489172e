#!/usr/bin/env python3
"""detect + describe of one large named workload on one GPU (default C5_blobs_1024), timed."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200.synth import workload, WORKLOADS
name = sys.argv[1] if len(sys.argv) > 1 else "C5_blobs_1024"
(nx, ny, nz), units, _, _ = WORKLOADS[name]
t0 = time.perf_counter()
im = s3.Image(nx, ny, nz, units=units)
workload(name, out=im.array)
t_syn = time.perf_counter() - t0
det = s3.Detector(0)
kp, ds = s3.KeypointStore(), s3.DescriptorStore()
out = {"workload": name, "synth_s": t_syn, "runs": []}
for rep in range(3):
    t0 = time.perf_counter()
    det.detect(im, kp)
    t1 = time.perf_counter()
    det.extract(kp, ds)
    t2 = time.perf_counter()
    out["runs"].append({"detect_ms": (t1 - t0) * 1e3, "describe_ms": (t2 - t1) * 1e3, "stage_ms": det.timings()})
out["candidates"] = len(det.candidates())
out["keypoints"] = len(kp)
out["voxels_per_s_e2e"] = nx * ny * nz / ((out["runs"][-1]["detect_ms"] + out["runs"][-1]["describe_ms"]) * 1e-3)
print(json.dumps(out))
