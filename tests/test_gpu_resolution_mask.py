"""Row a10: the resolution mask (apply_resolution_mask, spotfinder/kernels/masking.cu:37-73,99-147,
host caller spotfinder/spotfinder.cc:110-137) through the C ABI, against

  * the REFERENCE's own kernel, compiled in place for sm_100 (oracle/_ref/libffs_refgpu.so) and run
    on this GPU: must be identical, pixel for pixel (same float operations, same compiler);
  * the CPU restatement (oracle.resolution_mask, gcc libm, no FMA contraction): identical except
    pixels whose d lies within a few float ulps of dmin / dmax (device sqrtf/atanf/sinf and the
    fused dx*dx+dy*dy may round the last bit differently); that set must stay below 1e-5 of
    the pixels and every differing pixel must sit on the ring.
"""
import numpy as np
import pytest

from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle

pytestmark = pytest.mark.gpu

# (name, geometry, wavelength A, distance m, beam centre px, pixel size m)
DETECTORS = {
    "eiger16m": (synth.EIGER2_16M, 0.976, 0.306, (2074.3, 2181.7), 75e-6),
    "jungfrau4m": (synth.JUNGFRAU_4M, 1.2398, 0.120, (1030.2, 1100.9), 75e-6),
}
# dmin-only, dmax-only, both, beam centre off the detector
CASES = [
    ("dmin", 1.8, -1.0, None),
    ("dmax", -1.0, 20.0, None),
    ("both", 2.2, 12.0, None),
    ("off_detector", 1.5, 6.0, "off"),
]


def d_spacing_f64(g, wavelength, distance, bc, ps):
    y, x = np.mgrid[0:g.height, 0:g.width]
    dx = ((x + 0.5) - bc[0]) * ps
    dy = ((y + 0.5) - bc[1]) * ps
    r = np.sqrt(dx * dx + dy * dy)
    theta = 0.5 * np.arctan(r / distance)
    with np.errstate(divide="ignore"):
        return wavelength / (2 * np.sin(theta))


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("det", list(DETECTORS))
def test_resolution_mask_vs_reference_kernel_and_oracle(det, case):
    g, wavelength, distance, bc, ps = DETECTORS[det]
    _, dmin, dmax, bc_override = case
    if bc_override:  # beam centre beside the detector, the ring still crosses it
        bc = (-350.5, g.height + 300.25)
    mask = synth.make_mask(g)
    args = (np.float32(wavelength), np.float32(distance), np.float32(bc[0]), np.float32(bc[1]), np.float32(ps),
            np.float32(ps), np.float32(dmin), np.float32(dmax))
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=1) as sf:
        before = sf.mask_valid_count
        sf.apply_resolution_mask(*[float(a) for a in args])
        got = sf.get_mask()
        assert sf.mask_valid_count == int(got.sum())
    assert got.shape == mask.shape and set(np.unique(got)) <= {0, 1}
    assert not np.any(got[mask == 0]), "a masked pixel became valid"
    assert 0 < int(got.sum()) < before, "the bound removed nothing / everything"

    # 1. the reference's own kernel on this GPU: bit-exact
    if oracle.refgpu_available():
        ref = oracle.refgpu_resolution_mask(mask, *[float(a) for a in args])
        assert np.array_equal(got, ref), f"{int((got != ref).sum())} pixels differ from the reference kernel"

    # 2. the CPU restatement: equal off the ring
    want = oracle.resolution_mask(mask, *[float(a) for a in args])
    diff = got != want
    n_diff = int(diff.sum())
    assert n_diff < 1e-5 * g.pixels, f"{n_diff} pixels differ from the CPU oracle"
    if n_diff:
        d = d_spacing_f64(g, float(args[0]), float(args[1]), (float(args[2]), float(args[3])), float(args[4]))[diff]
        near = np.zeros(n_diff, bool)
        for bound in (dmin, dmax):
            if bound > 0:
                near |= np.abs(d - bound) <= 8 * np.spacing(np.float32(bound))
        assert near.all(), "a pixel away from the dmin/dmax ring differs"


def test_resolution_mask_feeds_the_threshold():
    """The calculated mask is what the threshold kernels then see (spotfinder.cc:650-652): strong
    pixels computed after ffs_apply_resolution_mask equal the oracle run with that mask."""
    g = synth.small_geometry(700, 500)
    mask = synth.make_mask(g)
    img = synth.make_frame(g, mask, np.uint16, config=7, frame_index=3, n_spots=60)
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=1) as sf:
        sf.apply_resolution_mask(1.0, 0.1, 350.0, 250.0, 75e-6, 75e-6, 4.0, 30.0)
        m2 = sf.get_mask()
        sf.upload(img)
        sf.threshold(0, 1)
        got = sf.strong_mask(0)
    assert 0 < m2.sum() < mask.sum()
    assert np.array_equal(got, oracle.dispersion(img, m2))
