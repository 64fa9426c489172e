"""Per-level blur times with and without an upload in flight (diagnostic)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sift3d_b200 as s3
from sift3d_b200 import synth
det = s3.Detector(0); det.set_profiling(True)
ims = []
for i in range(4):
    im = s3.Image(256, 256, 256); synth.workload("C2_mri_256", seed=100 + i, out=im.array); ims.append(im)
kp, ds = s3.KeypointStore(), s3.DescriptorStore()
def run(n, pre):
    acc = {}
    for i in range(n):
        if pre: ims[(i + 1) % 4].prefetch(0)
        det.detect(ims[i % 4], kp); det.extract(kp, ds)
        for o in range(2):
            for s in range(-1 if o == 0 else 0, 5):
                ms, _ = det.blur_ms(o, s)
                acc[(o, s)] = acc.get((o, s), 0.0) + ms / n
        t = det.timings()
        for k in ("scale", "pyramid", "extrema", "orient", "total_detect", "descriptors"):
            acc[k] = acc.get(k, 0.0) + t[k] / n
    return acc
for im in ims: im.make_resident(0)
run(4, False); a = run(20, False)
for im in ims: im.make_resident(-1)
ims[0].prefetch(0); run(4, True); b = run(20, True)
for k in a: print(k, f"{a[k]*1e3:8.1f} us   with upload in flight {b[k]*1e3:8.1f} us")
