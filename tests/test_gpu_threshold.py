"""Parity of the CUDA dispersion threshold (through the C ABI) against the CPU oracle:
bit-exact strong-pixel masks."""
import numpy as np
import pytest

from helpers import random_frame, unpack_bits
from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle

pytestmark = pytest.mark.gpu


def run_gpu(img, mask, **kw):
    h, w = img.shape
    with Spotfinder(w, h, img.dtype, mask, n_slots=1, **kw) as sf:
        sf.upload(img)
        sf.threshold(0, 1)
        u8 = sf.strong_mask(0)
        bits = unpack_bits(sf.strong_bits(0), w)
        assert np.array_equal(u8, bits)
        return u8


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("shape", [(64, 64), (33, 257), (100, 241), (7, 7), (1, 1), (3, 500), (500, 3),
                                   (129, 480), (240, 240), (131, 721)])
def test_dispersion_small(shape, dtype):
    h, w = shape
    for seed in range(3):
        img, mask = random_frame(h, w, dtype, seed=seed, lam=[0.2, 2.0, 20.0][seed])
        want = oracle.dispersion(img, mask)
        got = run_gpu(img, mask)
        assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_dispersion_random_mask_and_extremes(dtype):
    rng = np.random.default_rng(5)
    h, w = 150, 333
    img, mask = random_frame(h, w, dtype, seed=11, mask_kind="random")
    # saturated unmasked pixels, zeros, bright block
    img[rng.integers(0, h, 30), rng.integers(0, w, 30)] = np.iinfo(np.uint16).max
    img[40:48, 100:108] = 40000
    img[80:90, :] = 0
    if dtype == np.uint32:
        img[rng.integers(0, h, 10), rng.integers(0, w, 10)] = 2 ** 24  # excluded from sums
        img[rng.integers(0, h, 10), rng.integers(0, w, 10)] = 2 ** 24 - 1
        img[rng.integers(0, h, 5), rng.integers(0, w, 5)] = 2 ** 31
    want = oracle.dispersion(img, mask)
    got = run_gpu(img, mask)
    assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("lam", [9000.0, 9362.0, 30000.0])
@pytest.mark.parametrize("mask_kind", ["random", "modules"])
def test_dispersion_background_at_and_above_the_clamp(lam, mask_kind, dtype):
    """The streaming kernel's pruning bound is built from values clamped to 9362; a background
    at or above the clamp leaves (almost) nothing to prune.  Results must not change."""
    img, mask = random_frame(96, 300, dtype, seed=int(lam), lam=lam, n_spots=6, mask_kind=mask_kind)
    img[10:20, 30:60] = 9361
    img[30:40, 30:60] = 9362
    img[50:60, 30:60] = 9363
    want = oracle.dispersion(img, mask)
    got = run_gpu(img, mask)
    assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"


@pytest.mark.parametrize("lam", [0.02, 0.1, 0.5, 1.0])
@pytest.mark.parametrize("nsig_b", [6.0, 1.0, 0.0])
def test_dispersion_sparse_background(lam, nsig_b):
    """Background far below one count per pixel: nearly every non-zero pixel passes the signal
    bound and reaches the exact test; parity must hold for any nsig_b."""
    img, mask = random_frame(200, 520, np.uint16, seed=int(lam * 100) + 3, lam=lam, n_spots=25, mask_kind="modules")
    img[50:53, 100:140] = np.where(mask[50:53, 100:140] != 0, 1, img[50:53, 100:140])  # runs of single counts
    img[120, 200:230:2] = np.where(mask[120, 200:230:2] != 0, 3, img[120, 200:230:2])
    want = oracle.dispersion(img, mask, oracle.default_params(nsig_b=nsig_b))
    got = run_gpu(img, mask, nsig_b=nsig_b)
    assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"
    assert want.sum() > 0


@pytest.mark.parametrize("frac", [0.002, 0.02, 0.2])
@pytest.mark.parametrize("nsig", [(6.0, 3.0), (0.5, 0.5), (2.0, 8.0)])
def test_dispersion_small_pixel_exact_path_and_handover(frac, nsig):
    """16-bit full-window pixels take a 32-bit exact test when every window pixel is < 256
    (resolve_pixel_full_small) and are handed to the general routine otherwise: sprinkle values
    on both sides of 255/256 over a low background so that both paths, the handover queue (more
    than a warp's worth of handovers per flush at frac = 0.2) and many near-threshold decisions
    are exercised; all-valid mask so that the full-window path is the one taken."""
    nsig_b, nsig_s = nsig
    h, w = 260, 760
    rng = np.random.default_rng(int(frac * 1000) + int(nsig_b * 10))
    img = rng.poisson(2.0, (h, w)).astype(np.uint16)
    hot = rng.random((h, w)) < frac
    img[hot] = rng.choice(np.array([200, 254, 255, 256, 257, 300, 1000, 9362, 9363, 65535], np.uint16), int(hot.sum()))
    img[100:140, 300:340] = rng.poisson(180.0, (40, 40))  # a patch whose windows stay below 256 but are not small
    mask = np.ones((h, w), np.uint8)
    want = oracle.dispersion(img, mask, oracle.default_params(nsig_b=nsig_b, nsig_s=nsig_s))
    got = run_gpu(img, mask, nsig_b=nsig_b, nsig_s=nsig_s)
    assert np.array_equal(got, want), f"{np.argwhere(got != want)[:10]}"
    assert want.sum() > 0


def test_dispersion_all_zero_all_masked_no_mask():
    h, w = 60, 300
    z = np.zeros((h, w), np.uint16)
    ones = np.ones((h, w), np.uint8)
    assert run_gpu(z, ones).sum() == 0
    img, _ = random_frame(h, w, seed=2)
    assert run_gpu(img, np.zeros((h, w), np.uint8)).sum() == 0
    with Spotfinder(w, h, np.uint16, None, n_slots=1) as sf:  # NULL mask = all valid
        got = sf.standard_dispersion(img)
    assert np.array_equal(got, oracle.dispersion(img, ones).astype(bool))


def test_dispersion_params():
    img, mask = random_frame(200, 500, seed=4, lam=5.0, n_spots=20)
    for kw in (dict(nsig_b=3.0, nsig_s=1.5), dict(min_count=10), dict(threshold=5.0),
               dict(nsig_s=0.0), dict(min_count=0)):
        want = oracle.dispersion(img, mask, oracle.default_params(**kw))
        got = run_gpu(img, mask, **kw)
        assert np.array_equal(got, want), kw


@pytest.mark.parametrize("geom,dtype", [(synth.EIGER2_4M, np.uint16), (synth.JUNGFRAU_4M, np.uint32),
                                        (synth.EIGER2_16M, np.uint16)])
def test_dispersion_full_frames_vs_reference(geom, dtype):
    mask = synth.make_mask(geom)
    img = synth.make_frame(geom, mask, dtype, config=2, frame_index=1)
    assert synth.domain_ok(img, mask)
    want = oracle.reference_dispersion(img, mask) if oracle.ref_available() else oracle.dispersion(img, mask)
    got = run_gpu(img, mask)
    assert want.sum() > 100
    assert np.array_equal(got, want), f"{(got != want).sum()} differing pixels"


def test_batch_slots_independent():
    g = synth.small_geometry(700, 300)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 5, np.uint16, config=7)
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=5) as sf:
        sf.upload(frames)
        sf.threshold(0, 5)
        for i in range(5):
            assert np.array_equal(sf.strong_mask(i), oracle.dispersion(frames[i], mask)), i
        # re-run a sub-range: results unchanged, other slots untouched
        sf.threshold(1, 2)
        for i in range(5):
            assert np.array_equal(sf.strong_mask(i), oracle.dispersion(frames[i], mask)), i


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_candidate_overflow_and_brute_paths(dtype, monkeypatch):
    """A full candidate list must fall back to the exact all-pixel resolve, not drop pixels."""
    img, mask = random_frame(300, 700, dtype, seed=3, lam=3.0, n_spots=30)
    want = oracle.dispersion(img, mask)
    monkeypatch.setenv("FFS_CAND_CAP", "16")
    assert np.array_equal(run_gpu(img, mask), want)
    monkeypatch.delenv("FFS_CAND_CAP")
    monkeypatch.setenv("FFS_FORCE_BRUTE", "1")
    assert np.array_equal(run_gpu(img, mask), want)
