"""Shared test helpers: seeded inputs and oracle access (oracle/ is the checker only)."""
import numpy as np

from ffs_b200 import synth
from oracle import oracle


def random_frame(h, w, dtype=np.uint16, seed=0, lam=2.0, n_spots=8, mask_kind="modules"):
    g = synth.small_geometry(w, h)
    if mask_kind == "modules" and w > 24 and h > 24:
        mask = synth.make_mask(g, seed=seed, dead_fraction=2e-3)
    elif mask_kind == "random":
        rng = np.random.default_rng(seed)
        mask = (rng.random((h, w)) > 0.2).astype(np.uint8)
    else:
        mask = np.ones((h, w), np.uint8)
    img = synth.make_frame(g, mask, dtype, config=9, frame_index=seed, background=lam, n_spots=n_spots,
                           n_hot=3)
    return img, mask


def unpack_bits(bits, width):
    return np.unpackbits(bits, axis=1, bitorder="little")[:, :width]
