"""GPU parity tests proper: the CUDA path through the drop-in C API against the oracle
(oracle/_ref = the unmodified reference compiled from /root/reference; see
oracle/Makefile) on the same seeded inputs. Run on the B200 box with `-m gpu`."""
import numpy as np
import pytest

import parity
from sift3d_b200.synth import synth_volume

pytestmark = pytest.mark.gpu


def _ref():
    from oracle import refstage
    assert refstage.available(), "oracle/_ref/libsift3d_refstage.so missing (make -C oracle)"
    return refstage


def _gauss_taps(sigma):
    """init_Gauss_filter (imutil.c:1267-1319) in numpy, for the blur unit tests."""
    hw = max(int(np.ceil(sigma * 3.0)), 1)
    x = (np.arange(2 * hw + 1, dtype=np.float64) - hw) / (sigma + np.finfo(np.float64).eps)
    k = np.exp(-0.5 * x * x).astype(np.float32)
    acc = np.float32(0)
    for v in k:
        acc = np.float32(acc + v)
    return (k / acc).astype(np.float32)


@pytest.fixture(scope="module")
def det():
    import sift3d_b200 as s3
    d = s3.Detector()
    yield d
    d.close()


# Tolerance of the fast arithmetic mode on ONE blur of unit-variance noise: one FFMA per tap in the
# fused kernel (one rounding instead of two), banded multiply-adds for general spacings (the sum
# reordered): measured ~1e-7 of the level's max-abs; the contract (BASELINE.json north_star) is 1e-5.
FAST_BLUR_REL_TOL = 2e-6


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "fast"])
@pytest.mark.parametrize("units", [(1, 1, 1), (2, 2, 2), (4, 4, 4), (8, 8, 8), (0.75, 0.75, 1.5),
                                   (1.0, 2.0, 0.5)])
@pytest.mark.parametrize("sigma", [0.5387, 0.9733, 1.5450, 2.4525])
def test_blur_vs_reference(det, units, sigma, exact):
    """One separable blur against apply_Sep_FIR_filter incl. all boundaries, odd shapes (the
    literal kernels): exact mode bit for bit, fast mode within FAST_BLUR_REL_TOL."""
    ref = _ref()
    rng = np.random.default_rng(int(sigma * 1000) + int(units[0] * 8))
    vol = rng.standard_normal((23, 20, 37)).astype(np.float32)
    taps = _gauss_taps(sigma)
    want = ref.ref_blur(vol, units, taps)
    det.set_exact(exact)
    try:
        got = det.blur(vol, units, taps)
    finally:
        det.set_exact(False)
    rel, nbits = parity.level_error(got, want)
    if exact:
        assert nbits == 0, f"{nbits} voxels differ (max rel err {rel:.3g})"
    else:
        assert rel <= FAST_BLUR_REL_TOL, f"max rel err {rel:.3g}"


@pytest.mark.parametrize("dims", [(64, 48, 40), (96, 72, 45), (36, 32, 50), (128, 64, 37), (64, 48, 9)])
@pytest.mark.parametrize("units", [(2, 2, 2), (4, 4, 4), (8, 8, 8), (0.75, 0.75, 1.5), (1.0, 2.0, 0.5),
                                   (0.9, 0.9, 1.4), (1.3, 0.7, 1.0), (0.25, 0.3, 3.0)])
@pytest.mark.parametrize("sigma", [0.5387, 1.2263, 2.4525])
def test_banded_blur_vs_reference(det, dims, units, sigma):
    """Fast mode, general tap spacing (anisotropic units, octaves >= 2): the banded multiply-add
    kernels (blur_band.cuh; 3 launches) within FAST_BLUR_REL_TOL of apply_Sep_FIR_filter on all
    voxels incl. the six faces (mirror, -0.1 mirror, extrapolated samples, carried coordinates)."""
    ref = _ref()
    nx, ny, nz = dims
    rng = np.random.default_rng(int(sigma * 1000) + nx + int(units[0] * 16))
    vol = rng.standard_normal((nz, ny, nx)).astype(np.float32)
    taps = _gauss_taps(sigma)
    hw = len(taps) // 2
    if any(int(np.ceil(hw / u)) > n - 2 for u, n in zip(units, dims)):
        # a mirrored coordinate leaves the line: the reference reads outside it (imutil.c:833-854
        # mirrors once), there is nothing to compare with
        pytest.skip("filter reach longer than the line: undefined in the reference")
    want = ref.ref_blur(vol, units, taps)
    assert not det.exact()
    l0 = det.launch_count()
    got = det.blur(vol, units, taps)
    launches = det.launch_count() - l0
    rel, _ = parity.level_error(got, want)
    assert rel <= FAST_BLUR_REL_TOL, f"max rel err {rel:.3g}"
    assert launches == 3  # (lines shorter than their band still fit one: the band is then the line)


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "fast"])
@pytest.mark.parametrize("dims", [(64, 64, 40), (96, 70, 45), (36, 33, 50), (128, 64, 37),
                                  (32, 32, 36), (260, 100, 41), (64, 160, 150)])
@pytest.mark.parametrize("sigma", [0.3, 0.5387, 0.9733, 1.2263, 1.5450, 1.9466, 2.2, 2.4525])
def test_fused_blur_vs_reference(det, dims, sigma, exact):
    """The fused single-launch kernel (unit tap spacing, TMA-staged tiles, register z
    window) against apply_Sep_FIR_filter: half-widths 1..8, full / shifted / overlapping
    tiles, z-chunked and single-chunk volumes, all six boundary faces. Exact mode: bit for
    bit; fast mode (FFMA): within FAST_BLUR_REL_TOL of the level's max-abs."""
    ref = _ref()
    nx, ny, nz = dims
    rng = np.random.default_rng(int(sigma * 1000) + nx)
    vol = rng.standard_normal((nz, ny, nx)).astype(np.float32)
    taps = _gauss_taps(sigma)
    want = ref.ref_blur(vol, (1, 1, 1), taps)
    det.set_exact(exact)
    try:
        l0 = det.launch_count()
        got = det.blur(vol, (1, 1, 1), taps)
        # the fused TMA kernel really ran: one launch, not the three generic passes
        assert det.launch_count() - l0 == 1
        rel, nbits = parity.level_error(got, want)
        if exact:
            assert nbits == 0, f"{nbits} voxels differ (max rel err {rel:.3g})"
        else:
            assert rel <= FAST_BLUR_REL_TOL, f"max rel err {rel:.3g}"
        l0 = det.launch_count()
        gen = det.blur(vol, (1, 1, 1), taps, force_generic=True)
        assert det.launch_count() - l0 == 3
        assert parity.level_error(gen, want)[1] == 0
    finally:
        det.set_exact(False)


@pytest.mark.parametrize("dims", [(64, 64, 40), (96, 70, 45), (36, 33, 50), (128, 64, 37),
                                  (32, 32, 36), (128, 128, 128)])
@pytest.mark.parametrize("sigma", [0.3, 0.5387, 0.9733, 1.2263, 1.5450, 1.9466, 2.2, 2.4525])
def test_fused_half_blur_bit_exact(det, dims, sigma):
    """The fused kernel for tap spacing 1/2 (octave 1: units (2,2,2), unit 1) ==
    apply_Sep_FIR_filter bit for bit: half-widths 1..8 (odd and even), the -1/2 extrapolated
    sample, mirrored and "-0.1" virtual samples on all six faces, shifted tiles, z chunks."""
    ref = _ref()
    nx, ny, nz = dims
    rng = np.random.default_rng(int(sigma * 1000) + nx + 7)
    vol = rng.standard_normal((nz, ny, nx)).astype(np.float32)
    taps = _gauss_taps(sigma)
    want = ref.ref_blur(vol, (2, 2, 2), taps)
    det.set_exact(True)  # (the fast mode runs octave 1 on the banded passes: test_banded_blur_vs_reference)
    try:
        l0 = det.launch_count()
        got = det.blur(vol, (2, 2, 2), taps)
        assert det.launch_count() - l0 == 1  # the fused spacing-1/2 kernel, not the generic passes
        rel, nbits = parity.level_error(got, want)
        assert nbits == 0, f"{nbits} voxels differ (max rel err {rel:.3g})"
        gen = det.blur(vol, (2, 2, 2), taps, force_generic=True)
        assert parity.level_error(gen, want)[1] == 0
    finally:
        det.set_exact(False)


CASES = [
    # name, dims (nx, ny, nz), units, blobs, flags
    ("blobs_64", (64, 64, 64), (1.0, 1.0, 1.0), 64, 0),
    ("ragged_70x45x33", (70, 45, 33), (1.0, 1.0, 1.0), 40, 0),
    ("aniso_64x64x48", (64, 64, 48), (0.75, 0.75, 1.5), 48, 0),
    ("noisy_80", (80, 80, 80), (1.0, 1.0, 1.0), 125, 3),
]


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "fast"])
@pytest.mark.parametrize("name,dims,units,blobs,flags", CASES)
def test_full_pipeline_vs_reference(name, dims, units, blobs, flags, exact):
    _ref()
    vol = synth_volume(*dims, num_blobs=blobs, flags=flags)
    want = parity.run_reference(vol, units)
    got = parity.run_product(vol, units, exact=exact)
    parity.compare_runs(got, want, exact)
    iw, ig = zip(*parity.match_keypoints(want["keys"], got["keys"]))
    np.testing.assert_allclose(got["kp_matrix"][list(ig)], want["kp_matrix"][list(iw)],
                               atol=parity.KP_POS_TOL)


# Non-default detector parameters (setters sift.h:40-78) against the reference.
PARAM_CASES = [
    ("thresholds", dict(peak_thresh=0.05, corner_thresh=0.3)),
    ("levels_sigma", dict(num_kp_levels=2, sigma0=2.0, sigma_n=0.9)),
    ("four_levels", dict(num_kp_levels=4, peak_thresh=0.08)),
]


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "fast"])
@pytest.mark.parametrize("name,params", PARAM_CASES)
def test_non_default_parameters_vs_reference(name, params, exact):
    """peak_thresh / corner_thresh / num_kp_levels / sigma0 / sigma_n set through the public
    setters on both sides: same pyramid geometry, candidates, keypoints, descriptors."""
    _ref()
    vol = synth_volume(72, 64, 56, num_blobs=60, flags=3, seed=99)
    want = parity.run_reference(vol, (1.0, 1.0, 1.0), params)
    got = parity.run_product(vol, (1.0, 1.0, 1.0), params, exact=exact)
    parity.compare_runs(got, want, exact)


def test_cli_sequence_128_top100():
    """BASELINE.json configs[0]: kpSift3D's call sequence (cli/kpSift3D.c:116-135) on a
    128^3 blob volume: detect, keep the 100 'strongest' (with the reference's strength
    misalignment), describe."""
    _ref()
    vol = synth_volume(128, 128, 128)
    want = parity.run_reference(vol, limit=100)
    # exact arithmetic: the 'strongest 100' are picked by strengths that belong to OTHER
    # candidates (D8), so one flipped marginal extremum would shift every later slot
    got = parity.run_product(vol, limit=100, exact=True)
    assert parity.cand_tuple_list(got["candidates"]) == parity.cand_tuple_list(want["candidates"])
    assert len(got["keys"]) == len(want["keys"]) == 100
    pairs = parity.match_keypoints(want["keys"], got["keys"])
    assert len(pairs) >= 99
    # same order after the sort, same (misaligned) strengths
    same_slot = sum(1 for i, j in pairs if i == j)
    assert same_slot >= 99
    iw = [i for i, _ in pairs]
    ig = [j for _, j in pairs]
    np.testing.assert_array_equal(got["keys"]["strength"][ig], want["keys"]["strength"][iw])
    cos = parity.cosine_rows(got["desc"][ig, 3:], want["desc"][iw, 3:])
    assert cos.min() >= parity.DESC_COS_MIN


def test_api_error_behaviour():
    """Failure modes of the public API match the reference (SURVEY.md §8b 'Errors')."""
    import sift3d_b200 as s3
    d = s3.Detector()
    assert d.set_peak_thresh(0.0) == -1 and d.set_peak_thresh(1.5) == -1 and d.set_peak_thresh(1.0) == 0
    assert d.set_corner_thresh(-0.1) == -1 and d.set_corner_thresh(1.1) == -1 and d.set_corner_thresh(0.4) == 0
    assert d.set_sigma_n(-1.0) == -1 and d.set_sigma0(-1.0) == -1
    assert d.set_peak_thresh(0.1) == 0
    # extract before any detect
    kp = s3.KeypointStore()
    with pytest.raises(s3.Sift3DError):
        d.extract(kp)
    # image smaller than 8 voxels in a dimension (sift.c:447-452)
    with pytest.raises(s3.Sift3DError):
        d.detect(s3.Image.from_array(np.ones((7, 16, 16), np.float32)))
    # all-zero image: no keypoints, success; extract on an empty store fails (sift.c:1178)
    kp = d.detect(s3.Image.from_array(np.zeros((16, 16, 16), np.float32)))
    assert len(kp) == 0
    assert kp.to_matrix().shape == (0, 3)
    with pytest.raises(s3.Sift3DError):
        d.extract(kp)
    # multi-channel image rejected (sift.c:1221-1226)
    with pytest.raises(s3.Sift3DError):
        d.detect(s3.Image(16, 16, 16, 2))
    d.close()


def test_detector_reuse_and_replan():
    """Same detector, second image with different dims re-plans; same dims does not."""
    import sift3d_b200 as s3
    _ref()
    d = s3.Detector()
    for dims, blobs in (((48, 48, 48), 27), ((40, 56, 32), 20), ((40, 56, 32), 21)):
        vol = synth_volume(*dims, num_blobs=blobs, seed=1234 + blobs)
        want = parity.run_reference(vol)
        kp = d.detect(s3.Image.from_array(vol))
        got = d.candidates()
        assert parity.cand_tuple_list(got) == parity.cand_tuple_list(want["candidates"])
        keys = kp.records()
        assert len(parity.match_keypoints(want["keys"], keys)) >= 0.99 * len(want["keys"])
    d.close()


def test_detector_reuse_mixed_units():
    """Octaves >= 1 keep the units of the last re-plan (sift.c:651-656): dims X / units u1,
    then dims X / u2 (no re-plan: octaves >= 1 stay at u1 * 2^o), then dims Y / u2 (re-plan).
    Every step must match a reference detector driven through the same sequence, so cached
    per-plan state (orientation window tables) may not survive a change of units."""
    import sift3d_b200 as s3
    refstage = _ref()
    u1, u2 = (1.0, 1.0, 1.0), (0.8, 0.8, 1.6)
    seq = [((48, 48, 48), u1, 27), ((48, 48, 48), u2, 28), ((40, 56, 32), u2, 20)]
    d = s3.Detector()
    r = None
    for dims, units, blobs in seq:
        vol = synth_volume(*dims, num_blobs=blobs, seed=4321 + blobs)
        if r is None:
            r = refstage.RefRun(vol, units)
        else:
            r.set_image(vol, units)
        want_keys = r.detect()
        kp = d.detect(s3.Image.from_array(vol, units))
        assert parity.cand_tuple_list(d.candidates()) == parity.cand_tuple_list(r.candidates)
        keys = kp.records()
        pairs = parity.match_keypoints(want_keys, keys)
        assert len(pairs) >= 0.99 * len(want_keys), (dims, units, len(pairs), len(want_keys))
        iw = [i for i, _ in pairs]
        ig = [j for _, j in pairs]
        np.testing.assert_allclose(keys["R"][ig], want_keys["R"][iw], atol=2e-5)
    d.close()


def test_descriptor_cluster_split_is_bit_identical(monkeypatch):
    """Fewer keypoints than CTA slots: 2 or 4 CTAs (a thread-block cluster) share a keypoint and
    rank 0 adds their integer block histograms through distributed shared memory -- the
    descriptors must have the same bits as with one CTA per keypoint."""
    import sift3d_b200 as s3
    vol = s3.synth_volume(96, 80, 72, num_blobs=60, seed=4242)
    det = s3.Detector()
    kp = det.detect(s3.Image.from_array(vol, (1.0, 1.0, 1.0)))
    assert 8 <= len(kp) <= 296  # few enough for the 4-way split to be the default
    out = {}
    for split in ("1", "2", "4", None):
        if split is None:
            monkeypatch.delenv("SIFT3D_CUDA_DESC_SPLIT", raising=False)
        else:
            monkeypatch.setenv("SIFT3D_CUDA_DESC_SPLIT", split)
        out[split] = det.extract(kp).to_matrix().copy()
    det.close()
    for split in ("2", "4", None):
        assert np.array_equal(out[split].view(np.uint32), out["1"].view(np.uint32)), split
    assert np.abs(out["1"][:, 3:]).max() > 0
