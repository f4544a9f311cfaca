"""baseline.h-compatible C API (spotfinder_create / _standard_dispersion / _free) vs the
reference's own C API semantics: strong count returned, borrowed bool plane."""
import ctypes as C

import numpy as np
import pytest

from helpers import random_frame
from ffs_b200 import lib
from oracle import oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype,create,run", [(np.uint16, "spotfinder_create", "spotfinder_standard_dispersion"),
                                              (np.uint32, "spotfinder_create32", "spotfinder_standard_dispersion32")])
def test_baseline_c_api(dtype, create, run):
    L = lib.load()
    h, w = 120, 300
    ctx = getattr(L, create)(w, h)
    assert ctx
    try:
        for seed in range(3):  # the mask changes between calls, as the reference API allows
            img, mask = random_frame(h, w, dtype, seed=seed, mask_kind=["modules", "random", "ones"][seed])
            it = lib.ImageT(img.ctypes.data, mask.ctypes.data, h, w)
            dst = C.c_void_p()
            n = getattr(L, run)(ctx, C.byref(it), C.byref(dst))
            got = np.ctypeslib.as_array((C.c_uint8 * (h * w)).from_address(dst.value)).reshape(h, w)
            want = oracle.dispersion(img, mask)
            assert n == int(want.sum())
            assert np.array_equal(got, want)
    finally:
        L.spotfinder_free(ctx)
