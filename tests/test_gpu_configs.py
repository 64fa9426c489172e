"""GPU parity at BASELINE.json's own configurations: config 2 (the 256^3 MRI-like volume the
bench times) and config 3 (512 x 512 x 300, anisotropic units (0.75, 0.75, 1.5), descriptors
at all keypoints), CUDA path through the drop-in C API against the unmodified reference
(oracle/_ref) on the same seeded volume, in both blur arithmetic modes. The reference run of a
configuration (8 s / about a minute of host time) is shared by the two modes."""
import numpy as np
import pytest

import parity
from sift3d_b200.synth import workload

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(1500)]

_cache = {}


def _reference(name):
    from oracle import refstage
    assert refstage.available(), "oracle/_ref/libsift3d_refstage.so missing (make -C oracle)"
    if name not in _cache:
        _cache.clear()  # one configuration's reference pyramids at a time (config 3: ~8 GB)
        vol, units = workload(name)
        _cache[name] = (vol, units, parity.run_reference(vol, units))
    return _cache[name]


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "fast"])
@pytest.mark.parametrize("name", ["C2_mri_256", "C3_ct_512x512x300"])
def test_baseline_config_vs_reference(name, exact):
    vol, units, want = _reference(name)
    got = parity.run_product(vol, units, exact=exact)
    rep = parity.compare_runs(got, want, exact)
    print(name, "exact" if exact else "fast", rep)
    # the counts SURVEY.md §8(a) quotes for config 3 (measured on the reference)
    if name == "C3_ct_512x512x300":
        assert rep["candidates"][1] == 33588 and rep["keypoints"][1] == 9424
    got["det"].close()
