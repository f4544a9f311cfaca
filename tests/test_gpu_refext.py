"""The device threshold kernels (through the C ABI) against the REFERENCE'S OWN CPU classes of
baseline/spotfinder/baseline.cpp compiled in place (oracle/_ref/libffs_refext.so, stand-in scitbx /
DIALS headers: oracle/stub_dials): dispersion_extended stage by stage, and the gain-map variants
of both algorithms -- the same comparison tests/test_oracle_refext.py makes for the oracle, here
without the oracle in between.  SURVEY.md 8(a) rows a3, a4, a8.  Bit for bit."""
import numpy as np
import pytest

from helpers import random_frame
from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle
from test_gpu_extended import run_ext
from test_gpu_gain import ext_planes

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not oracle.refext_available(), reason="oracle/_ref/libffs_refext.so not built")]
D, X = oracle.ALGO_DISPERSION, oracle.ALGO_DISPERSION_EXTENDED


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_extended_three_stages_vs_reference_classes(dtype):
    for seed, (h, w, lam) in enumerate([(100, 241, 1.0), (129, 480, 4.0), (300, 700, 0.2)]):
        img, mask = random_frame(h, w, dtype, seed=seed, lam=lam, n_spots=25)
        s1, s2, strong, _ = run_ext(img, mask)
        assert np.array_equal(s1, oracle.refext_threshold(img, mask, X, stage=1)), "compute_dispersion_threshold"
        assert np.array_equal(s2, oracle.refext_threshold(img, mask, X, stage=2)), "erode_dispersion_mask"
        assert np.array_equal(strong, oracle.refext_threshold(img, mask, X)), "compute_final_threshold"


def test_extended_eiger4m_frame_vs_reference_classes():
    g = synth.EIGER2_4M
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 1, np.uint16, config=2, n_spots=300)
    s1, s2, strong, _ = run_ext(frames[0], mask)
    assert np.array_equal(strong, oracle.refext_threshold(frames[0], mask, X))
    assert np.array_equal(s1, oracle.refext_threshold(frames[0], mask, X, stage=1))
    assert strong.sum() > 1000


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_gain_maps_vs_reference_classes(dtype):
    h, w = 200, 520
    img, mask = random_frame(h, w, dtype, seed=33, lam=3.0, n_spots=30)
    rng = np.random.default_rng(4)
    modules = np.ones((h, w))
    modules[:, 130:260], modules[:, 260:390], modules[100:, 390:] = 1.07, 0.94, 1.21
    for gain in (modules, rng.uniform(0.25, 4.0, (h, w))):
        with Spotfinder(w, h, dtype, mask, n_slots=1) as sf:          # DispersionThreshold::threshold_w_gain
            sf.set_gain(gain)
            sf.upload(img)
            sf.threshold(0, 1)
            assert np.array_equal(sf.strong_mask(0), oracle.refext_threshold(img, mask, D, gain=gain))
        with Spotfinder(w, h, dtype, mask, n_slots=1, algorithm="dispersion_extended") as sf:
            sf.set_gain(gain)                                          # DispersionExtendedThreshold::threshold_w_gain
            got, (d1, b2) = ext_planes(sf, img)
            assert np.array_equal(d1, oracle.refext_threshold(img, mask, X, gain=gain, stage=1))
            assert np.array_equal(b2, oracle.refext_threshold(img, mask, X, gain=gain, stage=2))
            assert np.array_equal(got, oracle.refext_threshold(img, mask, X, gain=gain))
