"""Pins the oracle's connected components, spot fields, reflection filters, 3D components and
Kabsch-space variances (oracle/ffs_oracle.c, oracle.kabsch_variances) against the REFERENCE'S OWN
host code running here: spotfinder/connected_components/connected_components.cc and dx2's Panel /
Scan, compiled in place from /root/reference into oracle/_ref/libffs_refcc{,32}.so
(oracle/Makefile target `refcc`; Boost.Graph and Eigen are replaced by the stand-in headers of
oracle/stub_cc, see its README).  SURVEY.md 8(a) rows a12-a15 and 8(f) rank 3.

Everything integer is compared bit for bit, in the reference's own output order; centres of
mass are the reference's floats, bit for bit; variances to 1e-12 relative (the oracle restates
the same fp64 sequence with numpy)."""
import numpy as np
import pytest

from ffs_b200 import synth
from oracle import oracle

pytestmark = pytest.mark.skipif(not (oracle.refcc_available(2) and oracle.refcc_available(4)),
                                reason="oracle/_ref/libffs_refcc*.so not built (make -C oracle refcc; needs /root/reference)")

INT_FIELDS = ("x_min", "x_max", "y_min", "y_max", "z_min", "z_max", "num_pixels")


def assert_same_spots(ref, ours):
    assert len(ref) == len(ours)
    for name in INT_FIELDS:
        assert np.array_equal(ref[name].astype(np.int64), ours[name].astype(np.int64)), name
    for name in ("com_x", "com_y", "com_z"):
        assert np.array_equal(ref[name], ours[name].astype(np.float32)), name


def random_case(rng, h, w, density, dtype, plateau=False):
    strong = (rng.random((h, w)) < density).astype(np.uint8)
    hi = 8 if plateau else (60000 if dtype == np.uint16 else 3_000_000)
    image = rng.integers(1, hi, (h, w)).astype(dtype)
    return strong, image


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
@pytest.mark.parametrize("density", [0.02, 0.2, 0.55])
def test_2d_components_boxes_and_filters(dtype, density):
    rng = np.random.default_rng(int(density * 100) + np.dtype(dtype).itemsize)
    for h, w, min_size, max_sep, plateau in [(40, 64, 3, 2.0, False), (33, 47, 1, 0.7, True), (64, 31, 5, 1.3, True)]:
        strong, image = random_case(rng, h, w, density, dtype, plateau)
        _, spots = oracle.connected_components(strong, image, want_labels=False)
        with oracle.RefCC(image, strong, min_size) as r:
            assert r.num_strong[0] == int(strong.sum())
            big = spots[spots["num_pixels"] >= min_size]
            # generate_boxes: one box per component of at least min_spot_size pixels, in label order
            assert r.num_strong_filtered[0] == int(big["num_pixels"].sum())
            b = r.boxes[0]
            assert len(b) == len(big)
            assert np.array_equal(b["l"], big["x_min"]) and np.array_equal(b["r"], big["x_max"])
            assert np.array_equal(b["t"], big["y_min"]) and np.array_equal(b["b"], big["y_max"])
            assert np.array_equal(b["num_pixels"], big["num_pixels"])
            # find_2d_components: both filters; ties of the peak intensity resolved as is_signal_preferred does
            ref, before = r.find_2d(0, min_size, max_sep)
            kept, rs, rp = oracle.filter_spots(spots, min_size, max_sep)
            assert before == len(big) and len(spots) - rs == len(big)
            assert_same_spots(ref, kept)


def test_row_wrap_quirk_is_the_reference_behaviour():
    """build_graph links a pixel to linear_index + 1 without looking at the row (connected_components.cc:
    62-70): the last pixel of a row and the first of the next are one component."""
    strong = np.zeros((6, 8), np.uint8)
    strong[1, 7] = strong[2, 0] = 1          # wrap pair
    strong[4, 3] = strong[4, 4] = strong[5, 4] = 1
    image = np.full((6, 8), 10, np.uint16)
    _, wrap = oracle.connected_components(strong, image, row_wrap=True, want_labels=False)
    _, nowrap = oracle.connected_components(strong, image, row_wrap=False, want_labels=False)
    with oracle.RefCC(image, strong, 1) as r:
        ref, before = r.find_2d(0, 1, 100.0)
    assert before == 2 and len(wrap) == 2 and len(nowrap) == 3
    assert_same_spots(ref, wrap)
    assert (ref["x_min"][0], ref["x_max"][0], ref["num_pixels"][0]) == (0, 7, 2)


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_3d_components_order_fields_and_text(dtype):
    g = synth.small_geometry(240, 180)
    mask = synth.make_mask(g)
    sweep = synth.make_sweep(g, mask, 9, dtype, n_reflections=50)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    labels, spots = oracle.connected_components(strong, sweep)
    for min_size, max_sep in [(3, 2.0), (1, 0.9), (6, 5.0)]:
        kept, rs, rp = oracle.filter_spots(spots, min_size, max_sep)
        with oracle.RefCC(sweep, strong, 3) as r:
            ref = r.find_3d(min_size, max_sep)
        assert len(ref) == len(spots) - rs - rp
        assert_same_spots(ref, kept)
        # 3d_reflections.txt lines (spotfinder.cc:1119-1148) in the reference's order
        assert oracle.format_reflections_3d(ref) == oracle.format_reflections_3d(kept)


def test_3d_random_volumes():
    rng = np.random.default_rng(11)
    for z, h, w, density in [(5, 20, 24, 0.08), (7, 16, 16, 0.3), (3, 30, 12, 0.5), (1, 10, 10, 0.3)]:
        strong = (rng.random((z, h, w)) < density).astype(np.uint8)
        image = rng.integers(1, 12, (z, h, w)).astype(np.uint16)     # many intensity ties
        _, spots = oracle.connected_components(strong, image, want_labels=False)
        kept, _, _ = oracle.filter_spots(spots, 2, 1.5)
        with oracle.RefCC(image, strong, 2) as r:
            assert r.num_strong == [int(s.sum()) for s in strong]
            ref = r.find_3d(2, 1.5)
        assert_same_spots(ref, kept)


def test_kabsch_variances_are_the_reference_numbers():
    g = synth.small_geometry(240, 180)
    mask = synth.make_mask(g)
    sweep = synth.make_sweep(g, mask, 10, np.uint16, n_reflections=60)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    labels, spots = oracle.connected_components(strong, sweep)
    kept, _, _ = oracle.filter_spots(spots, 3, 2.0)
    for geom in [dict(distance_mm=150.0, beam_center_x_px=120.0, beam_center_y_px=90.0, pixel_size_x_mm=0.075,
                      pixel_size_y_mm=0.075, wavelength=0.976, oscillation_width_deg=0.1),
                 dict(distance_mm=289.3, beam_center_x_px=31.7, beam_center_y_px=160.2, pixel_size_x_mm=0.172,
                      pixel_size_y_mm=0.172, wavelength=1.54, oscillation_width_deg=0.25)]:
        with oracle.RefCC(sweep, strong, 3) as r:
            r.find_3d(3, 2.0)
            ref = r.variances(geom, oscillation_start=12.5)   # the start angle cancels in eps3
        sb, sm, depth = oracle.kabsch_variances(labels, sweep, kept, geom)
        assert len(ref) == len(kept) > 20
        assert np.array_equal(ref["depth"], depth)
        assert np.allclose(ref["sigma_b_variance"], sb, rtol=1e-12, atol=0)
        assert np.allclose(ref["sigma_m_variance"], sm, rtol=1e-12, atol=1e-300)
        assert (ref["depth"] >= 5).any() and (ref["sigma_m_variance"] > 0).any()
