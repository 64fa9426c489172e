#!/usr/bin/env python3
"""Timeline of one detect: begin/end of every level's blur relative to the start of detect
(octave streams), plus the stage marks. Diagnostic for the schedule."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200 import api
from sift3d_b200.synth import workload
vol, units = workload(sys.argv[1] if len(sys.argv) > 1 else "C2_mri_256")
det = s3.Detector(0); det.set_profiling(True)
im = s3.Image.from_array(vol, units); im.make_resident(0)
kp = s3.KeypointStore()
for _ in range(3): det.detect(im, kp)
L = api.libcuda()
L.sift3d_cuda_level_span_ms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
for o in range(det.num_octaves()):
    row = []
    for s in range(-1 if o == 0 else 0, 5):
        a, b = C.c_float(), C.c_float()
        if L.sift3d_cuda_level_span_ms(det.ctx, o, s, C.byref(a), C.byref(b)) == 0:
            row.append(f"s{s}:{a.value*1e3:6.0f}-{b.value*1e3:6.0f}")
    print(f"octave {o}: " + "  ".join(row))
t = det.timings()
acc = 0.0
for k in ("h2d", "scale", "pyramid", "extrema", "order", "orient", "d2h"):
    acc += t[k]; print(f"{k:8s} ends at {acc*1e3:7.0f} us")
