"""Debug: compare fused-kernel strong mask with the oracle and print the differing pixels."""
import sys, os
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, "fast-feedback-service_b200"); sys.path.insert(0, ".")
from helpers import random_frame
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle

for (h, w) in [(64, 64), (100, 241), (300, 700)]:
    for dtype in (np.uint16, np.uint32):
        img, mask = random_frame(h, w, dtype, seed=0, lam=1.0, n_spots=25)
        want = oracle.dispersion(img, mask)
        with Spotfinder(w, h, img.dtype, mask, n_slots=1) as sf:
            sf.upload(img); sf.process(0, 1)
            got = sf.strong_mask(0)
        d = np.argwhere(got != want)
        print(h, w, dtype.__name__, "want", int(want.sum()), "got", int(got.sum()), "diff", len(d))
        for (y, x) in d[:12]:
            print("   y", y, "x", x, "got", got[y, x], "want", want[y, x], "v", img[y, x], "mask", mask[y, x])
