"""Pins the oracle's dispersion_extended and gain-map restatements (oracle/ffs_oracle.c) against the
REFERENCE'S OWN CPU classes running here: baseline/spotfinder/baseline.cpp (DispersionThreshold,
DispersionExtendedThreshold, each with threshold() and threshold_w_gain()), compiled in place
from /root/reference into oracle/_ref/libffs_refext.so (oracle/Makefile target `refext`; the scitbx
array views and DIALS's chebyshev_distance are the stand-ins of oracle/stub_dials, see its
README).  SURVEY.md 8(a) rows a3 and a4; the three stages of the extended algorithm are compared
one by one.  Every comparison is bit for bit."""
import numpy as np
import pytest

from helpers import random_frame
from ffs_b200 import synth
from oracle import oracle

pytestmark = pytest.mark.skipif(not oracle.refext_available(),
                                reason="oracle/_ref/libffs_refext.so not built (make -C oracle refext; needs /root/reference)")
D, X = oracle.ALGO_DISPERSION, oracle.ALGO_DISPERSION_EXTENDED


def same(a, b):
    return np.array_equal(np.asarray(a).astype(np.uint8), np.asarray(b).astype(np.uint8))


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_extended_stage_by_stage(dtype):
    for seed, (h, w, lam) in enumerate([(64, 64, 0.3), (131, 257, 2.0), (200, 300, 25.0), (37, 500, 0.05)]):
        img, mask = random_frame(h, w, dtype, seed=seed, lam=lam, n_spots=20)
        for stage in (1, 2, 3):
            want = oracle.refext_threshold(img, mask, X, stage=stage)
            assert same(oracle.extended_stage(img, mask, stage), want), (seed, stage)
        assert same(oracle.dispersion_extended(img, mask), oracle.refext_threshold(img, mask, X))
        assert same(oracle.dispersion(img, mask), oracle.refext_threshold(img, mask, D))


def test_extended_on_detector_geometry_with_module_gaps():
    g = synth.small_geometry(520, 400)
    mask = synth.make_mask(g)
    assert (mask == 0).any()
    for config in (1, 2, 8):
        frames = synth.make_frames(g, mask, 2, np.uint16, config=config, n_spots=40)
        for f in frames:
            want = oracle.refext_threshold(f, mask, X)
            assert same(oracle.dispersion_extended(f, mask), want)
            assert want.sum() > 0


def test_extended_parameters_and_edge_inputs():
    rng = np.random.default_rng(4)
    img, mask = random_frame(90, 120, np.uint32, seed=9, lam=3.0, n_spots=15)
    # pixels at and above 2^24 leave the sums (compute_sat BIG, baseline.cpp:378,386) but stay candidates
    img2 = img.copy()
    img2[rng.integers(0, 90, 40), rng.integers(0, 120, 40)] = np.array([1 << 24, (1 << 24) - 1, (1 << 24) + 5, 4_000_000_000] * 10, np.uint32)
    for p in [oracle.default_params(min_count=2), oracle.default_params(min_count=10), oracle.default_params(min_count=49),
              oracle.default_params(nsig_b=3.0, nsig_s=1.5), oracle.default_params(nsig_b=0.0, nsig_s=0.0),
              oracle.default_params(threshold=7.0), oracle.default_params(kernel_half_x=2, kernel_half_y=4, min_count=3),
              oracle.default_params(kernel_half_x=5, kernel_half_y=1, min_count=2)]:
        for im in (img, img2):
            for stage in (1, 2, 3):
                assert same(oracle.extended_stage(im, mask, stage, p), oracle.refext_threshold(im, mask, X, p, stage=stage))
            assert same(oracle.dispersion(im, mask, p), oracle.refext_threshold(im, mask, D, p))
    # all-zero frame, everything masked, nothing masked, a frame smaller than the kernel
    z = np.zeros((40, 50), np.uint16)
    ones = np.ones((40, 50), np.uint8)
    for im, m in [(z, ones), (img[:40, :50].astype(np.uint16), np.zeros((40, 50), np.uint8)),
                  (img[:40, :50].astype(np.uint16), ones), (img[:5, :4].astype(np.uint16), np.ones((5, 4), np.uint8))]:
        assert same(oracle.dispersion_extended(im, m), oracle.refext_threshold(im, m, X))
        assert same(oracle.dispersion(im, m), oracle.refext_threshold(im, m, D))


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_gain_variants(dtype):
    """threshold_w_gain of both classes (baseline.cpp:187-250,289-310 and :484-541,657-714,772-805)."""
    rng = np.random.default_rng(8)
    for seed, lam in enumerate([0.5, 4.0, 40.0]):
        img, mask = random_frame(110, 150, dtype, seed=seed + 20, lam=lam, n_spots=20)
        gains = [np.ones(img.shape), rng.uniform(0.5, 2.0, img.shape),
                 np.where(np.arange(img.shape[1])[None, :] < 75, 0.8, 1.3) * np.ones(img.shape)]
        for g in gains:
            assert same(oracle.dispersion_gain(img, mask, g), oracle.refext_threshold(img, mask, D, gain=g))
            for stage in (1, 2, 3):
                assert same(oracle.dispersion_extended_gain(img, mask, g, stage=stage),
                            oracle.refext_threshold(img, mask, X, gain=g, stage=stage)), (seed, stage)


def test_the_references_argument_checks():
    img, mask = random_frame(30, 30, np.uint16, seed=1, lam=1.0)
    with pytest.raises(ValueError):      # DIALS_ASSERT(min_count_ <= num_kernel && min_count_ > 1)
        oracle.refext_threshold(img, mask, X, oracle.default_params(min_count=1))
    with pytest.raises(ValueError):
        oracle.refext_threshold(img, mask, X, oracle.default_params(min_count=50))
