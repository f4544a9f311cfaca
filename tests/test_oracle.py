"""CPU-only: pins the oracle (oracle/ffs_oracle.c) against the reference's own CPU code
(golden vectors produced by standalone.cc, and the live oracle/_ref build when present),
and cross-checks the parts of the oracle that cannot be pinned offline."""
from pathlib import Path

import numpy as np
import pytest
from scipy import ndimage

from helpers import random_frame, unpack_bits
from oracle import oracle

GOLDEN = np.load(Path(__file__).parent / "golden" / "dispersion_ref.npz")
N_GOLDEN = sum(1 for k in GOLDEN.files if k.startswith("image_"))


@pytest.mark.parametrize("i", range(N_GOLDEN))
def test_oracle_matches_reference_golden(i):
    img, mask = GOLDEN[f"image_{i}"], GOLDEN[f"mask_{i}"]
    want = unpack_bits(GOLDEN[f"strong_{i}"], img.shape[1])
    assert np.array_equal(oracle.dispersion(img, mask), want)
    assert np.array_equal(oracle.dispersion_bruteforce(img, mask), want)


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_oracle_matches_live_reference(dtype):
    for seed, (h, w) in enumerate([(150, 333), (64, 64), (33, 500)]):
        img, mask = random_frame(h, w, dtype, seed=seed, lam=[0.2, 2.0, 20.0][seed], n_spots=20)
        assert np.array_equal(oracle.dispersion(img, mask), oracle.reference_dispersion(img, mask))


def test_extended_self_checks():
    """dispersion_extended is unpinned offline (baseline.cpp needs DIALS headers): the SAT
    restatement must equal the brute-force one, and its first pass AND-ed with the signal
    test must reproduce the (pinned) standard dispersion."""
    for seed, (h, w) in enumerate([(90, 140), (64, 200)]):
        img, mask = random_frame(h, w, np.uint16, seed=seed, lam=1.5, n_spots=15)
        ext = oracle.dispersion_extended(img, mask)
        assert np.array_equal(ext, oracle.dispersion_extended_bruteforce(img, mask))
        assert ext.sum() > 0
        s1 = oracle.extended_stage(img, mask, 1)
        s2 = oracle.extended_stage(img, mask, 2)
        std = oracle.dispersion(img, mask)
        assert not (std & ~s1).any()  # standard = first pass AND (v > threshold) AND (b > d)
        assert not (s2 & ~mask.astype(bool)).any()  # eroded plane is 0 on masked pixels
        assert not (ext & s2).any()  # strong pixels lie in the eroded signal core only


def test_connected_components_vs_scipy():
    """Boost.Graph is absent offline: check the union-find oracle against scipy's labelling
    on row-wrap-free input (true 4-connectivity), then the 3D (6-connectivity) case."""
    rng = np.random.default_rng(0)
    strong = (rng.random((60, 80)) > 0.7).astype(np.uint8)
    strong[:, -1] = 0  # no pixel in the last column: the row-wrap quirk cannot fire
    img = rng.integers(1, 1000, strong.shape).astype(np.uint16)
    labels, spots = oracle.connected_components(strong, img, row_wrap=True)
    ref, n = ndimage.label(strong, structure=ndimage.generate_binary_structure(2, 1))
    assert len(spots) == n
    # same partition: a bijection between label sets
    pairs = set(zip(labels[strong > 0].tolist(), ref[strong > 0].tolist()))
    assert len(pairs) == n
    assert np.array_equal(np.sort(spots["num_pixels"]), np.sort(np.bincount(ref.ravel())[1:]))
    # canonical label = minimum linear index, ascending order
    assert np.all(np.diff(spots["label"].astype(np.int64)) > 0)
    s3 = (rng.random((5, 30, 40)) > 0.75).astype(np.uint8)
    s3[:, :, -1] = 0
    i3 = rng.integers(1, 1000, s3.shape).astype(np.uint16)
    _, sp3 = oracle.connected_components(s3, i3, row_wrap=True, want_labels=False)
    _, n3 = ndimage.label(s3, structure=ndimage.generate_binary_structure(3, 1))
    assert len(sp3) == n3


def test_spot_fields_and_filters():
    strong = np.zeros((10, 12), np.uint8)
    img = np.zeros((10, 12), np.uint16)
    strong[2, 3:6] = 1
    img[2, 3:6] = [10, 30, 10]
    strong[3, 4] = 1
    img[3, 4] = 30  # tie with (4,2): the smaller (y, x) wins the peak
    strong[7, 7] = 1
    img[7, 7] = 5
    _, s = oracle.connected_components(strong, img)
    assert list(s["num_pixels"]) == [4, 1]
    a = s[0]
    assert (a["x_min"], a["x_max"], a["y_min"], a["y_max"]) == (3, 5, 2, 3)
    assert (a["peak_x"], a["peak_y"], a["peak_intensity"]) == (4, 2, 30)
    np.testing.assert_allclose(a["com_x"], (3.5 * 10 + 4.5 * 30 + 5.5 * 10 + 4.5 * 30) / 80, rtol=1e-6)
    np.testing.assert_allclose(a["com_y"], (2.5 * 50 + 3.5 * 30) / 80, rtol=1e-6)
    assert a["com_z"] == 0.5
    kept, rs, rp = oracle.filter_spots(s, 3, 2.0)
    assert (len(kept), rs, rp) == (1, 1, 0)


def test_dispersion_gain_against_direct_window_sums():
    """threshold_w_gain restatement (SAT, C) vs direct 7x7 integer window sums + the same fp64
    formulas in numpy (baseline.cpp:240-246); the gain = 1 map differs from the plain test only
    in where x*(m-1) enters (a vs c), so the two agree on ordinary data."""
    from ffs_b200 import synth
    g = synth.small_geometry(90, 70)
    mask = synth.make_mask(g)
    img = synth.make_frame(g, mask, np.uint16, 2, 3, n_spots=8, background=3.0)
    rng = np.random.default_rng(9)
    gain = rng.uniform(0.3, 3.0, img.shape)
    h, w = img.shape
    want = np.zeros((h, w), np.uint8)
    v = img.astype(np.int64)
    ok = (mask != 0) & (v < 2 ** 24)
    for j in range(h):
        for i in range(w):
            ys, xs = slice(max(j - 3, 0), j + 4), slice(max(i - 3, 0), i + 4)
            sel = ok[ys, xs]
            m = float(sel.sum())
            x = float(v[ys, xs][sel].sum())
            y = float((v[ys, xs][sel] ** 2).sum())
            if mask[j, i] and m >= 2 and img[j, i] > 0:
                a = m * y - x * x
                b = m * float(img[j, i]) - x
                c = gain[j, i] * x * (m - 1 + 6.0 * np.sqrt(2 * (m - 1)))
                d = 3.0 * np.sqrt(gain[j, i] * x * m)
                want[j, i] = a > c and b > d
    got = oracle.dispersion_gain(img, mask, gain)
    assert np.array_equal(got, want)
    assert np.array_equal(oracle.dispersion_gain(img, mask, np.ones(img.shape)), oracle.dispersion(img, mask))


def test_extended_gain_against_direct_window_sums():
    """DispersionExtendedThreshold::threshold_w_gain restatement (baseline.cpp:484-541, 657-714,
    772-805; parity unpinned) vs direct integer window sums and the same fp64 formulas in numpy:
    first pass on the 7x7 window of mask-valid pixels, final pass on the 11x11 window of the
    pixels the (gain-independent) erosion left valid."""
    from ffs_b200 import synth
    g = synth.small_geometry(80, 64)
    mask = synth.make_mask(g)
    img = synth.make_frame(g, mask, np.uint16, 2, 5, n_spots=8, background=3.0)
    rng = np.random.default_rng(11)
    gain = rng.uniform(0.3, 3.0, img.shape)
    h, w = img.shape
    v = img.astype(np.int64)
    ok = (mask != 0) & (v < 2 ** 24)
    want1 = np.zeros((h, w), np.uint8)
    for j in range(h):
        for i in range(w):
            ys, xs = slice(max(j - 3, 0), j + 4), slice(max(i - 3, 0), i + 4)
            sel = ok[ys, xs]
            m = float(sel.sum())
            x = float(v[ys, xs][sel].sum())
            y = float((v[ys, xs][sel] ** 2).sum())
            if mask[j, i] and m >= 2:
                want1[j, i] = (m * y - x * x) > gain[j, i] * x * (m - 1 + 6.0 * np.sqrt(2 * (m - 1)))
    got1 = oracle.dispersion_extended_gain(img, mask, gain, stage=1)
    assert np.array_equal(got1, want1)
    eroded = oracle.dispersion_extended_gain(img, mask, gain, stage=2)
    ok2 = (eroded != 0) & (v < 2 ** 24)
    want = np.zeros((h, w), np.uint8)
    for j in range(h):
        for i in range(w):
            ys, xs = slice(max(j - 5, 0), j + 6), slice(max(i - 5, 0), i + 6)
            sel = ok2[ys, xs]
            m = float(sel.sum())
            x = float(v[ys, xs][sel].sum())
            if mask[j, i] and not eroded[j, i] and img[j, i] > 0:
                mean = x / m if m >= 2 else 0.0
                want[j, i] = float(img[j, i]) >= mean + 3.0 * np.sqrt(gain[j, i] * mean)
    got = oracle.dispersion_extended_gain(img, mask, gain)
    assert np.array_equal(got, want)
    assert want.sum() > 0
    # a unit gain map and the plain algorithm agree on ordinary data (they differ only in rounding)
    assert np.array_equal(oracle.dispersion_extended_gain(img, mask, np.ones(img.shape)),
                          oracle.dispersion_extended(img, mask))
