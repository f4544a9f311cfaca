"""ffs_get_strong_u8_static: the allocation-free variant of ffs_get_strong_u8 (scratch in a file-scope
device array, csrc/api_noalloc.inl).  No counterpart in the reference.
Through the C ABI, against the existing entry point, bit-exact (a 0/1 plane), on
two calls in a row whose inputs differ -- the array keeps the previous call's contents, so a stale
result would show.  Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

from ffs_b200 import lib, synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle

pytestmark = pytest.mark.gpu


def static_u8(sf, slot):
    out = np.full((sf.height, sf.width), 7, np.uint8)
    rc = sf._L.ffs_get_strong_u8_static(sf._ctx, slot, out.ctypes.data)
    return rc, out


def test_static_scratch_matches_allocating_entry_point_twice_in_a_row():
    L = lib.load()
    g = synth.EIGER2_4M  # 4,485,690 pixels: just below the 4.5 Mi maximum
    assert g.pixels <= L.ffs_noalloc_max_pixels()
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 2, np.uint16, config=1)
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=2) as sf:
        sf.upload(frames)
        sf.threshold(0, 2)
        planes = []
        for slot in (0, 1):  # two calls in a row, different contents
            rc, got = static_u8(sf, slot)
            assert rc == 0
            want = sf.strong_mask(slot)  # ffs_get_strong_u8, the allocating entry point
            assert np.array_equal(got, want)
            assert np.array_equal(got, oracle.dispersion(frames[slot], mask))
            planes.append(got)
        assert not np.array_equal(planes[0], planes[1])
        rc, again = static_u8(sf, 0)  # and back: nothing of slot 1 is left in the scratch
        assert rc == 0 and np.array_equal(again, planes[0])
    # a smaller context on the same GPU reuses the same array
    h, w = 123, 457
    g2 = synth.small_geometry(w, h)
    m2 = synth.make_mask(g2)
    f2 = synth.make_frame(g2, m2, np.uint32, config=3, frame_index=1, n_spots=20)
    with Spotfinder(w, h, np.uint32, m2, n_slots=1) as sf:
        sf.upload(f2)
        sf.threshold(0, 1)
        rc, got = static_u8(sf, 0)
        assert rc == 0 and np.array_equal(got, sf.strong_mask(0))


def test_static_scratch_rejects_frames_beyond_the_maximum():
    g = synth.EIGER2_16M  # 18 M pixels > FFS_NOALLOC_MAX_PIXELS
    with Spotfinder(g.width, g.height, np.uint16, None, n_slots=1) as sf:
        out = np.zeros((g.height, g.width), np.uint8)
        rc = sf._L.ffs_get_strong_u8_static(sf._ctx, 0, out.ctypes.data)
        assert rc == -4  # FFS_ERR_CAPACITY
        msg = (sf._L.ffs_last_error(sf._ctx) or b"").decode()
        assert "FFS_NOALLOC_MAX_PIXELS" in msg and str(g.pixels) in msg
        assert not out.any()
