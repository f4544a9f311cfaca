"""Regenerates tests/golden/dispersion_ref.npz from the REFERENCE's own CPU spotfinder
(baseline/spotfinder/standalone.cc compiled in place -> oracle/_ref/libffs_ref.so).
Run in the build container (needs /root/reference):  python tests/golden/make_golden.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path[:0] = [str(ROOT), str(ROOT / "fast-feedback-service_b200"), str(ROOT / "tests")]
from helpers import random_frame  # noqa: E402
from oracle import oracle  # noqa: E402

CASES = [  # (h, w, dtype, seed, lam, mask_kind)
    (96, 128, "uint16", 1, 2.0, "modules"), (61, 250, "uint16", 2, 0.3, "random"),
    (120, 90, "uint32", 3, 20.0, "modules"), (40, 300, "uint32", 4, 2.0, "ones"),
    (7, 9, "uint16", 5, 5.0, "ones"),
]


def main():
    assert oracle.ref_available(), "oracle/_ref/libffs_ref.so missing: run make -C oracle"
    out = {}
    for i, (h, w, dt, seed, lam, mk) in enumerate(CASES):
        img, mask = random_frame(h, w, np.dtype(dt), seed=seed, lam=lam, mask_kind=mk)
        if dt == "uint32":  # exercise the >= 2^24 exclusion of the reference (standalone.cc:78,90)
            img[h // 2, w // 2] = 2 ** 24
            img[h // 3, w // 3] = 2 ** 24 + 5
        out[f"image_{i}"], out[f"mask_{i}"] = img, mask
        out[f"strong_{i}"] = np.packbits(oracle.reference_dispersion(img, mask), axis=1, bitorder="little")
    np.savez_compressed(Path(__file__).with_name("dispersion_ref.npz"), **out)
    print("wrote", len(CASES), "cases")


if __name__ == "__main__":
    main()
