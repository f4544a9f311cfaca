"""The device connected components, spot reduction, filters, 3D components and Kabsch variances
(through the C ABI) against the REFERENCE'S OWN host code compiled in place
(oracle/_ref/libffs_refcc*.so = spotfinder/connected_components/connected_components.cc + dx2 Panel /
Scan behind stand-in Boost.Graph / Eigen headers, oracle/stub_cc) -- the same comparison
tests/test_oracle_refcc.py makes for the oracle, here without the oracle in between.
SURVEY.md 8(a) rows a12-a15, 8(f) rank 3."""
import numpy as np
import pytest

from helpers import random_frame
from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle
from test_gpu_kabsch import GEOM, run_sweep

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not (oracle.refcc_available(2) and oracle.refcc_available(4)),
                                 reason="oracle/_ref/libffs_refcc*.so not built")]

INT_FIELDS = ("x_min", "x_max", "y_min", "y_max", "z_min", "z_max", "num_pixels")


def same_spots(ref, got):
    assert len(ref) == len(got)
    for f in INT_FIELDS:
        assert np.array_equal(ref[f].astype(np.int64), got[f].astype(np.int64)), f
    for f in ("com_x", "com_y", "com_z"):
        np.testing.assert_allclose(got[f], ref[f], rtol=1e-5, atol=0, err_msg=f)   # north_star tolerance; equal in practice


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_frame_counts_boxes_and_2d_reflections(dtype):
    for seed, (h, w) in enumerate([(100, 241), (300, 700)]):
        img, mask = random_frame(h, w, dtype, seed=seed, lam=1.5, n_spots=30)
        with Spotfinder(w, h, dtype, mask, n_slots=1) as sf:
            sf.upload(img)
            sf.process(0, 1)
            res, spots = sf.fetch(0)
            strong = sf.strong_mask(0)
            kept, _, _ = sf.filter_spots(spots, 3, 2.0)      # ffs_filter_spots, the product's filter_reflections
        with oracle.RefCC(img, strong, 3) as r:
            assert res.num_strong_pixels == r.num_strong[0]
            assert res.num_strong_pixels_filtered == r.num_strong_filtered[0]      # get_num_strong_pixels_filtered()
            assert res.n_spots_total == len(r.boxes[0])                            # get_boxes().size()
            big = spots[spots["num_pixels"] >= 3]
            assert np.array_equal(r.boxes[0]["l"], big["x_min"]) and np.array_equal(r.boxes[0]["b"], big["y_max"])
            ref, before = r.find_2d(0, 3, 2.0)
        assert res.n_spots_filtered == len(ref)
        same_spots(ref, kept)
        assert len(ref) > 3


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_sweep_3d_reflections_and_variances(dtype):
    g = synth.small_geometry(240, 280)
    mask = synth.make_mask(g)
    sweep = synth.make_sweep(g, mask, 12, dtype, n_reflections=70)
    with Spotfinder(g.width, g.height, dtype, mask, n_slots=2) as sf:
        strong = []
        sf.sweep_begin(len(sweep))
        for z in range(len(sweep)):
            sf.upload(sweep[z], 0)
            sf.process(0, 1)
            strong.append(sf.strong_mask(0))
            sf.sweep_record_slot(0, z)
        spots = sf.sweep_finish()
        sums, moments = sf.sweep_export_sums(), sf.sweep_export_moments()
        vb, vm, depth = sf.kabsch_variances(GEOM, spots, sums, moments)
        ours_kept, _, _ = sf.filter_spots(spots, 3, 2.0)
    strong = np.stack(strong)
    with oracle.RefCC(sweep, strong, 3) as r:
        everything = r.find_3d(1, 1e9)          # nothing filtered: every component, reference order
        ref_var = r.variances(GEOM)
        kept = r.find_3d(3, 2.0)
    same_spots(everything, spots)
    assert np.array_equal(ref_var["depth"], depth)
    np.testing.assert_allclose(vb, ref_var["sigma_b_variance"], rtol=1e-6, atol=1e-18)
    np.testing.assert_allclose(vm, ref_var["sigma_m_variance"], rtol=1e-6, atol=1e-18)
    same_spots(kept, ours_kept)
    assert len(kept) > 30 and (depth > 1).sum() > 5
