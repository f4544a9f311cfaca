"""threshold_w_gain (baseline.cpp:187-250,289-310): the standard dispersion test
with a per-pixel gain map, CUDA through the C ABI vs the CPU oracle, bit-exact."""
import numpy as np
import pytest

from helpers import random_frame
from ffs_b200 import synth
from ffs_b200.spotfinder import FfsError, Spotfinder
from oracle import oracle

pytestmark = pytest.mark.gpu


def gpu_mask(sf, img):
    sf.upload(img)
    sf.threshold(0, 1)
    return sf.strong_mask(0)


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("kind", ["ones", "smooth", "random"])
def test_gain_map_matches_oracle(kind, dtype):
    h, w = 180, 500
    img, mask = random_frame(h, w, dtype, seed=21, lam=3.0, n_spots=25)
    rng = np.random.default_rng(3)
    if kind == "ones":
        gain = np.ones((h, w))
    elif kind == "smooth":
        yy, xx = np.mgrid[0:h, 0:w]
        gain = 0.6 + 1.2 * (xx / w) + 0.3 * np.sin(yy / 17.0)
    else:
        gain = rng.uniform(0.25, 4.0, (h, w))
    want = oracle.dispersion_gain(img, mask, gain)
    with Spotfinder(w, h, dtype, mask, n_slots=1) as sf:
        plain = gpu_mask(sf, img)
        sf.set_gain(gain)
        got = gpu_mask(sf, img)
        assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"
        sf.set_gain(None)  # back to the plain test
        assert np.array_equal(gpu_mask(sf, img), plain)
    assert np.array_equal(plain, oracle.dispersion(img, mask))
    assert want.sum() > 0
    if kind == "random":
        assert (want != plain).any()


def test_gain_with_nonpositive_entries_and_params():
    """A non-positive gain makes sqrt(gain*x*m) NaN in the reference: those pixels fail b > d;
    the kernel must not prune with such a map."""
    h, w = 120, 300
    img, mask = random_frame(h, w, np.uint16, seed=5, lam=2.0, n_spots=15)
    gain = np.full((h, w), 1.5)
    gain[40:60, 100:150] = 0.0
    gain[70:80, 10:60] = -1.0
    p = oracle.default_params(nsig_b=4.0, nsig_s=2.0, threshold=1.0)
    want = oracle.dispersion_gain(img, mask, gain, p)
    with Spotfinder(w, h, np.uint16, mask, n_slots=1, nsig_b=4.0, nsig_s=2.0, threshold=1.0) as sf:
        sf.set_gain(gain)
        got = gpu_mask(sf, img)
    assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"


def test_gain_streaming_after_graph_capture():
    g = synth.small_geometry(320, 200)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 6, np.uint16, config=6, n_spots=12)
    gain = np.random.default_rng(1).uniform(0.5, 2.0, mask.shape)
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=2) as sf:
        for i in range(4):  # slot graphs get captured without a gain map
            sf.submit_copy(i, frames[i])
            sf.wait()
        sf.set_gain(gain)
        for i in range(6):
            sf.submit_copy(i, frames[i])
            res, _ = sf.wait()
            assert res.num_strong_pixels == int(oracle.dispersion_gain(frames[i], mask, gain).sum())


def ext_planes(sf, img):
    sf.upload(img)
    sf.threshold(0, 1)
    return sf.strong_mask(0), (sf.extended_stage(0, 1), sf.extended_stage(0, 2))


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("kind", ["ones", "modules", "random"])
def test_extended_gain_map_matches_oracle(kind, dtype):
    """DispersionExtendedThreshold::threshold_w_gain (baseline.cpp:484-541,657-714,772-805): the
    streaming first pass brackets gain*x*K(m) by the smallest / largest gain and decides the
    pixels in between exactly; the final pass reads the gain per core pixel."""
    h, w = 200, 520
    img, mask = random_frame(h, w, dtype, seed=33, lam=3.0, n_spots=30)
    rng = np.random.default_rng(4)
    if kind == "ones":
        gain = np.ones((h, w))
    elif kind == "modules":  # per-module gains, as a real detector calibration would give
        gain = np.ones((h, w))
        gain[:, 130:260] = 1.07
        gain[:, 260:390] = 0.94
        gain[100:, 390:] = 1.21
    else:
        gain = rng.uniform(0.25, 4.0, (h, w))
    want = oracle.dispersion_extended_gain(img, mask, gain)
    want1 = oracle.dispersion_extended_gain(img, mask, gain, stage=1)
    want2 = oracle.dispersion_extended_gain(img, mask, gain, stage=2)
    with Spotfinder(w, h, dtype, mask, n_slots=1, algorithm="dispersion_extended") as sf:
        plain, _ = ext_planes(sf, img)
        sf.set_gain(gain)
        got, (d1, b2) = ext_planes(sf, img)
        assert np.array_equal(d1, want1), f"pass 1: {np.argwhere(d1 != want1)[:10]}"
        assert np.array_equal(b2, want2), f"erosion: {np.argwhere(b2 != want2)[:10]}"
        assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"
        sf.set_gain(None)
        assert np.array_equal(ext_planes(sf, img)[0], plain)
    assert np.array_equal(plain, oracle.dispersion_extended(img, mask))
    assert want.sum() > 0
    if kind == "random":
        assert (want1 != oracle.extended_stage(img, mask, 1)).any()


def test_extended_gain_with_nonpositive_entries():
    """gain <= 0: c <= 0 in the first pass (so a > c holds almost everywhere there) and
    sqrt(gain*mean) is NaN in the final pass; no bracket exists, every pixel is decided exactly."""
    h, w = 130, 310
    img, mask = random_frame(h, w, np.uint16, seed=8, lam=2.0, n_spots=15)
    gain = np.full((h, w), 1.3)
    gain[30:70, 90:160] = 0.0
    gain[80:95, 10:80] = -2.0
    p = oracle.default_params(nsig_b=4.0, nsig_s=2.0, threshold=1.0)
    want = oracle.dispersion_extended_gain(img, mask, gain, p)
    with Spotfinder(w, h, np.uint16, mask, n_slots=1, algorithm="dispersion_extended", nsig_b=4.0, nsig_s=2.0,
                    threshold=1.0) as sf:
        sf.set_gain(gain)
        got, _ = ext_planes(sf, img)
    assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"
