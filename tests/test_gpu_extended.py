"""dispersion_extended (PIXEL_DATA_32BIT and 16-bit) through the C ABI vs the CPU oracle,
stage by stage: first-pass mask, eroded plane, final strong mask; then CC on top."""
import numpy as np
import pytest

from helpers import random_frame
from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle
from test_gpu_cc import check_spots

pytestmark = pytest.mark.gpu


def run_ext(img, mask, **kw):
    h, w = img.shape
    with Spotfinder(w, h, img.dtype, mask, algorithm="dispersion_extended", n_slots=1, **kw) as sf:
        sf.upload(img)
        sf.process(0, 1)
        return sf.extended_stage(0, 1), sf.extended_stage(0, 2), sf.strong_mask(0), sf.fetch(0)


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("shape", [(64, 64), (100, 241), (37, 300), (5, 5), (129, 480)])
def test_extended_stages(shape, dtype):
    h, w = shape
    for seed in range(2):
        img, mask = random_frame(h, w, dtype, seed=seed, lam=1.0 + 3 * seed, n_spots=20)
        s1, s2, strong, (res, spots) = run_ext(img, mask)
        assert np.array_equal(s1, oracle.extended_stage(img, mask, 1)), "first pass"
        assert np.array_equal(s2, oracle.extended_stage(img, mask, 2)), "eroded plane"
        want = oracle.dispersion_extended(img, mask)
        assert np.array_equal(strong, want), "final"
        _, want_spots = oracle.connected_components(want, img, want_labels=False)
        assert res.num_strong_pixels == int(want.sum())
        check_spots(spots, want_spots)


def test_extended_blob_touching_gap_and_edges():
    """Masked pixels erode adjacent signal (baseline.cpp:558-569), image edges do not."""
    h, w = 90, 200
    rng = np.random.default_rng(2)
    img = rng.poisson(2.0, (h, w)).astype(np.uint32)
    mask = np.ones((h, w), np.uint8)
    mask[:, 100:106] = 0
    img[:, 100:106] = 2 ** 32 - 1
    img[20:32, 88:100] += 400      # blob touching the gap
    img[0:9, 0:9] += 400           # blob in the image corner
    img[50:62, 150:162] += 400     # free-standing blob
    s1, s2, strong, _ = run_ext(img, mask)
    assert np.array_equal(s1, oracle.extended_stage(img, mask, 1))
    assert np.array_equal(s2, oracle.extended_stage(img, mask, 2))
    want = oracle.dispersion_extended(img, mask)
    assert want.sum() > 50
    assert np.array_equal(strong, want)


def test_extended_params_and_overflow(monkeypatch):
    img, mask = random_frame(150, 400, np.uint16, seed=7, lam=4.0, n_spots=30)
    for kw in (dict(nsig_b=3.0, nsig_s=1.0), dict(min_count=20), dict(threshold=6.0)):
        want = oracle.dispersion_extended(img, mask, oracle.default_params(**kw))
        assert np.array_equal(run_ext(img, mask, **kw)[2], want), kw
    monkeypatch.setenv("FFS_CAND_CAP", "8")  # core list overflows -> all-pixel final pass
    assert np.array_equal(run_ext(img, mask)[2], oracle.dispersion_extended(img, mask))


def test_extended_eiger16m_32bit_module_gaps():
    """BASELINE.json configs[2]: Eiger2 16M 32-bit, extended two-pass, module-gap mask."""
    g = synth.EIGER2_16M
    mask = synth.make_mask(g)
    img = synth.make_frame(g, mask, np.uint32, config=3, frame_index=0)
    want = oracle.dispersion_extended(img, mask)
    _, _, strong, (res, _) = run_ext(img, mask)
    assert want.sum() > 100
    assert np.array_equal(strong, want)
    assert res.num_strong_pixels == int(want.sum())
