"""Trusted range (reader.get_trusted_range()[1] -> max_valid_pixel_value, spotfinder/spotfinder.cc:
482,868,879; candidate rejection spotfinder/kernels/thresholding.cu:208-209,317).  Decision SURVEY N3:
a mask-valid pixel ABOVE trusted_max is invalid for the candidate AND for its neighbours' window
sums, identically in oracle and kernel -- i.e. the result equals the CPU algorithm run with
mask' = mask && v <= trusted_max.  The kernel gets the ORIGINAL mask and the trusted_max
parameter; the oracle gets mask'.  Bit-exact, both algorithms, both pixel widths, the extended
algorithm's intermediate planes, counters and connected components included."""
import numpy as np
import pytest

from helpers import random_frame
from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle

pytestmark = pytest.mark.gpu


def overexposed(h, w, dtype, seed, tmax, n_bad=25, lam=2.0):
    img, mask = random_frame(h, w, dtype, seed=seed, lam=lam, n_spots=30)
    rng = np.random.default_rng(100 + seed)
    ys, xs = rng.integers(0, h, n_bad), rng.integers(0, w, n_bad)
    top = min(np.iinfo(dtype).max, 2 ** 24 + 5) + 1  # exclusive
    img[ys, xs] = rng.integers(tmax + 1, top, n_bad).astype(dtype)
    img[ys[0], xs[0]] = tmax + 1            # just above: invalid (tmax < type max in every case here)
    img[ys[1], xs[1]] = tmax                # exactly the cutoff: still valid (v <= max_valid, thresholding.cu:209)
    # a cluster: overlapping neighbourhoods, and a bright spot right next to an invalid pixel
    cy, cx = h // 2, w // 3
    img[cy:cy + 2, cx:cx + 3] = min(tmax + 7, int(np.iinfo(dtype).max))
    img[cy + 2, cx:cx + 3] = np.minimum(tmax, 900)
    if mask[0, 0]:
        img[0, 0] = tmax + 1                # image corner
    return img, mask


@pytest.mark.parametrize("algorithm", ["dispersion", "dispersion_extended"])
@pytest.mark.parametrize("dtype,tmax", [(np.uint16, 30000), (np.uint16, 65534), (np.uint32, 126634),
                                        (np.uint32, 2 ** 24 + 1), (np.uint16, 500)])
def test_trusted_range_equals_oracle_on_folded_mask(dtype, tmax, algorithm):
    fn = oracle.dispersion if algorithm == "dispersion" else oracle.dispersion_extended
    for seed, (h, w) in enumerate([(180, 500), (97, 263)]):
        img, mask = overexposed(h, w, dtype, seed, tmax)
        folded = (mask.astype(bool) & (img <= tmax)).astype(np.uint8)
        assert folded.sum() < mask.sum()
        want = fn(img, folded)
        assert not np.array_equal(want, fn(img, mask)), "the case must be sensitive to the trusted range"
        with Spotfinder(w, h, dtype, mask, n_slots=1, algorithm=algorithm, trusted_max=tmax) as sf:
            sf.upload(img)
            sf.process(0, 1)
            got = sf.strong_mask(0)
            assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"
            if algorithm == "dispersion_extended":
                assert np.array_equal(sf.extended_stage(0, 1), oracle.extended_stage(img, folded, 1))
                assert np.array_equal(sf.extended_stage(0, 2), oracle.extended_stage(img, folded, 2))
            res, spots = sf.fetch(0)
            _, spots_want = oracle.connected_components(want, img, want_labels=False)
            assert res.num_strong_pixels == int(want.sum())
            assert res.n_components == len(spots_want)
            assert np.array_equal(spots["label"], spots_want["label"])
            assert np.array_equal(spots["num_pixels"], spots_want["num_pixels"])
            assert np.array_equal(spots["sum_intensity"], spots_want["sum_intensity"])


def test_trusted_range_off_and_type_max_change_nothing():
    img, mask = overexposed(150, 300, np.uint16, 3, 40000)
    want = oracle.dispersion(img, mask)
    for tmax in (0, -1, 65535, 10 ** 9):
        with Spotfinder(300, 150, np.uint16, mask, n_slots=1, trusted_max=tmax) as sf:
            sf.upload(img)
            sf.threshold(0, 1)
            assert np.array_equal(sf.strong_mask(0), want)


def test_trusted_range_full_frame_batch():
    """Eiger2 4M-sized batch: every frame of a batch is repaired independently, repeatedly."""
    g = synth.EIGER2_4M
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 3, np.uint16, config=1)  # 20 unmasked hot pixels at 60000 per frame
    tmax = 50000
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=3, trusted_max=tmax) as sf:
        sf.upload(frames)
        for _ in range(2):
            sf.process(0, 3)
            for i in range(3):
                folded = (mask.astype(bool) & (frames[i] <= tmax)).astype(np.uint8)
                want = oracle.dispersion(frames[i], folded)
                assert np.array_equal(sf.strong_mask(i), want)
                res, _ = sf.fetch(i, 16)
                assert res.num_strong_pixels == int(want.sum())
