"""Parity of device connected components + spot reduction (through the C ABI) against the
CPU oracle: labels, pixel counts, bboxes, peaks bit-exact; centroids within 1e-5 relative
(they come out identical: the sums are exact integers on both sides)."""
import numpy as np
import pytest

from helpers import random_frame
from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle

pytestmark = pytest.mark.gpu

INT_FIELDS = ["label", "x_min", "x_max", "y_min", "y_max", "z_min", "z_max", "num_pixels", "peak_x",
              "peak_y", "peak_z", "peak_intensity", "sum_intensity"]
FLT_FIELDS = ["com_x", "com_y", "com_z", "peak_centroid_distance"]


def check_spots(got, want):
    assert len(got) == len(want)
    for f in INT_FIELDS:
        assert np.array_equal(got[f], want[f]), f
    for f in FLT_FIELDS:
        np.testing.assert_allclose(got[f], want[f], rtol=1e-5, atol=0, err_msg=f)


def run_case(img, mask, row_wrap=True, **kw):
    h, w = img.shape
    want_strong = oracle.dispersion(img, mask)
    want_labels, want_spots = oracle.connected_components(want_strong, img, row_wrap=row_wrap)
    with Spotfinder(w, h, img.dtype, mask, n_slots=1, cc_row_wrap=int(row_wrap), **kw) as sf:
        sf.upload(img)
        sf.process(0, 1)
        res, spots = sf.fetch(0)
        labels = sf.labels(0)
        assert np.array_equal(sf.strong_mask(0), want_strong)
    assert res.status == 0
    assert res.num_strong_pixels == int(want_strong.sum())
    assert res.n_components == len(want_spots)
    assert np.array_equal(labels, want_labels)
    check_spots(spots, want_spots)
    kept, rs, rp = oracle.filter_spots(want_spots, kw.get("min_spot_size", 3),
                                       kw.get("max_peak_centroid_separation", 2.0))
    boxes = want_spots[want_spots["num_pixels"] >= kw.get("min_spot_size", 3)]
    assert res.n_spots_total == len(boxes)
    assert res.num_strong_pixels_filtered == int(boxes["num_pixels"].sum())
    assert res.n_spots_filtered == len(kept)
    return res, spots


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("shape", [(64, 64), (100, 241), (300, 700), (131, 1000)])
def test_cc_small(shape, dtype):
    h, w = shape
    for seed in range(2):
        img, mask = random_frame(h, w, dtype, seed=seed, lam=1.0 + seed, n_spots=25)
        res, _ = run_case(img, mask)
        assert res.n_components > 3


@pytest.mark.parametrize("row_wrap", [True, False])
def test_cc_row_wrap_quirk(row_wrap):
    """connected_components.cc:62-70 links index k and k+1 across the row end."""
    h, w = 80, 96
    rng = np.random.default_rng(1)
    img = rng.poisson(1.0, (h, w)).astype(np.uint16)
    mask = np.ones((h, w), np.uint8)
    for y in (10, 30, 50):
        img[y, w - 2:] = 300      # right edge of row y
        img[y + 1, :2] = 300      # left edge of row y+1
    strong = oracle.dispersion(img, mask)
    assert strong[10, w - 1] and strong[11, 0]
    _, sp_wrap = oracle.connected_components(strong, img, row_wrap=True)
    _, sp_nowrap = oracle.connected_components(strong, img, row_wrap=False)
    assert len(sp_wrap) < len(sp_nowrap)  # the quirk really merges components here
    run_case(img, mask, row_wrap=row_wrap)


def test_cc_filters_and_empty():
    img, mask = random_frame(200, 400, seed=5, lam=0.5, n_spots=40)
    run_case(img, mask, min_spot_size=1, max_peak_centroid_separation=0.5)
    run_case(img, mask, min_spot_size=6, max_peak_centroid_separation=10.0)
    z = np.zeros((50, 300), np.uint16)
    with Spotfinder(300, 50, np.uint16, None, n_slots=1) as sf:
        sf.upload(z)
        sf.process(0, 1)
        res, spots = sf.fetch(0)
    assert res.num_strong_pixels == 0 and res.n_components == 0 and len(spots) == 0


def test_cc_large_blob_and_capacity():
    """One big component (deep union-find chains) and the capacity status."""
    h, w = 256, 512
    rng = np.random.default_rng(3)
    img = rng.poisson(1.0, (h, w)).astype(np.uint16)
    img[40:44, 5:500] = 800         # long horizontal bar
    img[40:200, 250:254] = 900      # vertical bar crossing it
    mask = np.ones((h, w), np.uint8)
    res, spots = run_case(img, mask)
    assert spots["num_pixels"].max() > 800
    with Spotfinder(w, h, np.uint16, mask, n_slots=1, max_strong_per_frame=64) as sf:
        sf.upload(img)
        sf.process(0, 1)
        res, _ = sf.fetch(0)
    assert res.status == -4  # FFS_ERR_CAPACITY, never a silent truncation


@pytest.mark.parametrize("geom,dtype", [(synth.EIGER2_16M, np.uint16), (synth.JUNGFRAU_4M, np.uint32)])
def test_cc_full_frames_batch(geom, dtype):
    mask = synth.make_mask(geom)
    frames = synth.make_frames(geom, mask, 3, dtype, config=2)
    with Spotfinder(geom.width, geom.height, dtype, mask, n_slots=3) as sf:
        sf.upload(frames)
        sf.process(0, 3)
        for i in range(3):
            strong = oracle.dispersion(frames[i], mask)
            _, want = oracle.connected_components(strong, frames[i], want_labels=False)
            res, spots = sf.fetch(i)
            assert res.num_strong_pixels == int(strong.sum())
            check_spots(spots, want)
            counts = oracle.process_frame(frames[i], mask)
            assert res.n_spots_total == counts["n_spots_total"]
            assert res.num_strong_pixels_filtered == counts["num_strong_pixels_filtered"]
            assert res.n_spots_filtered == counts["n_spots_filtered"]
