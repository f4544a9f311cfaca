"""Per-spot Kabsch-space variances: device second-moment reduction + closed
form (cc3_moments_kernel, kabsch_kernel) against the per-pixel fp64 restatement of
Reflection3D::variances_in_kabsch_space (oracle.kabsch_variances), through the C ABI."""
import numpy as np
import pytest

from ffs_b200 import synth
from ffs_b200.lib import SPOT_DTYPE
from ffs_b200.spotfinder import Spotfinder
from ffs_b200.sweep import merge_slabs, slab_bounds, slab_result_from
from oracle import oracle

pytestmark = pytest.mark.gpu

GEOM = dict(distance_mm=306.765, beam_center_x_px=119.7, beam_center_y_px=141.2, pixel_size_x_mm=0.075,
            pixel_size_y_mm=0.075, wavelength=0.976254, oscillation_width_deg=0.1)
RTOL = 1e-6  # fp64 closed form vs fp64 per-pixel sum: rounding only (north_star: floats within 1e-5)


def run_sweep(sf, sweep, z0=0, z1=None):
    z1 = len(sweep) if z1 is None else z1
    sf.sweep_begin(z1 - z0)
    for z in range(z0, z1):
        sf.upload(sweep[z], 0)
        sf.process(0, 1)
        sf.sweep_record_slot(0, z)


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_kabsch_variances_match_per_pixel_oracle(dtype):
    g = synth.small_geometry(240, 280)
    mask = synth.make_mask(g)
    sweep = synth.make_sweep(g, mask, 12, dtype, n_reflections=70)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    labels, want_spots = oracle.connected_components(strong, sweep)
    want = oracle.kabsch_variances(labels, sweep, want_spots, GEOM)
    with Spotfinder(g.width, g.height, dtype, mask, n_slots=2) as sf:
        run_sweep(sf, sweep)
        spots = sf.sweep_finish()
        sums, moments = sf.sweep_export_sums(), sf.sweep_export_moments()
        got = sf.kabsch_variances(GEOM, spots, sums, moments)
    assert np.array_equal(spots["label"], want_spots["label"])
    assert len(spots) > 30 and (got[2] > 1).sum() > 5  # really three-dimensional spots
    assert np.array_equal(got[2], want[2])
    for k, name in ((0, "sigma_b_variance"), (1, "sigma_m_variance")):
        np.testing.assert_allclose(got[k], want[k], rtol=RTOL, atol=1e-18, err_msg=name)


def test_kabsch_variances_after_slab_merge():
    """Moments exported per slab and merged on the host give the same variances as one sweep."""
    g = synth.small_geometry(200, 160)
    mask = synth.make_mask(g)
    sweep = synth.make_sweep(g, mask, 10, np.uint16, n_reflections=50)
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=2) as sf:
        run_sweep(sf, sweep)
        spots = sf.sweep_finish()
        whole = sf.kabsch_variances(GEOM, spots, sf.sweep_export_sums(), sf.sweep_export_moments())
        slabs = []
        for r in range(3):
            z0, z1 = slab_bounds(len(sweep), 3, r)
            run_sweep(sf, sweep, z0, z1)
            slabs.append(slab_result_from(sf, z0, z1 - z0))
        m_spots, m_sums, m_moments = merge_slabs(slabs, g.width, g.height, with_sums=True)
        merged = sf.kabsch_variances(GEOM, m_spots, m_sums, m_moments)
    assert np.array_equal(m_spots["label"], spots["label"])
    assert np.array_equal(merged[2], whole[2])
    np.testing.assert_allclose(merged[0], whole[0], rtol=1e-9, atol=1e-20)
    np.testing.assert_allclose(merged[1], whole[1], rtol=1e-9, atol=1e-20)


def test_kabsch_empty_and_zero_intensity():
    with Spotfinder(64, 64, np.uint16, None, n_slots=1) as sf:
        sb, sm, d = sf.kabsch_variances(GEOM, np.zeros(0, SPOT_DTYPE), np.zeros((0, 4), np.uint64),
                                        np.zeros((0, 4), np.uint64))
        assert len(sb) == len(sm) == len(d) == 0
