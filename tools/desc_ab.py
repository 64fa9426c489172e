#!/usr/bin/env python
"""A/B of two descriptor kernels on the bench workload: run in two processes (the kernel is
chosen by environment at first use) and compare the descriptor matrices.
    python tools/desc_ab.py run out.npy      (honours SIFT3D_CUDA_DESC_V3 / _ROWS)
    python tools/desc_ab.py cmp a.npy b.npy"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if sys.argv[1] == "oracle":
    # descriptor error of the active kernel against the CPU oracle on a small volume
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
    import parity
    from sift3d_b200 import synth
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 96
    vol = synth.synth_volume(n, n, n, seed=77, flags=3)
    got, want = parity.run_product(vol), parity.run_reference(vol)
    a, b = got["desc"], want["desc"]
    assert a.shape == b.shape, (a.shape, b.shape)
    da = np.abs(a[:, 3:] - b[:, 3:])
    cos = parity.cosine_rows(a[:, 3:], b[:, 3:])
    print("vs oracle: rows", a.shape[0], "max abs diff", da.max(), "mean abs diff", da.mean(),
          "min cosine", cos.min(), "rows with diff > 1e-5:", int((da.max(1) > 1e-5).sum()))
elif sys.argv[1] == "run":
    from sift3d_b200 import api, synth
    name = sys.argv[3] if len(sys.argv) > 3 else "C2_mri_256"
    vol, units = synth.workload(name)
    det = api.Detector(0)
    img = api.Image.from_array(vol, units)
    kp = det.detect(img)
    d = det.extract(kp).to_matrix()
    for _ in range(3):
        d = det.extract(kp).to_matrix()
    print(len(kp), det.timings()["descriptors"])
    np.save(sys.argv[2], d)
else:
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    assert a.shape == b.shape, (a.shape, b.shape)
    da = np.abs(a[:, 3:] - b[:, 3:])
    cos = (a[:, 3:] * b[:, 3:]).sum(1) / np.linalg.norm(a[:, 3:], axis=1) / np.linalg.norm(b[:, 3:], axis=1)
    print("rows", a.shape[0], "max abs diff", da.max(), "mean abs diff", da.mean(), "min cosine", cos.min(),
          "rows with diff > 1e-5:", int((da.max(1) > 1e-5).sum()))
