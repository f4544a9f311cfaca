"""Deterministic synthetic diffraction frames.

Poisson background + Gaussian spots + hot pixels on detector geometries with
module-gap masks.  Gap pixels carry the pixel type's maximum value, as real
Eiger data does, so any leak of masked pixels into a window sum shows up at once.

This module is workload generation only (numpy); it is not part of the hot path.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

BASE_SEED = 20261104


@dataclass(frozen=True)
class Geometry:
    name: str
    width: int
    height: int
    mod_fast: int
    mod_slow: int
    n_fast: int
    n_slow: int
    gap_fast: int
    gap_slow: int

    @property
    def pixels(self) -> int:
        return self.width * self.height


# reference: h5read/include/eiger2xe.h:6-17 (Eiger2 XE 16M module layout)
EIGER2_16M = Geometry("eiger2_16m", 4148, 4362, 1028, 512, 4, 8, 12, 38)
# BASELINE.json config 1 "Eiger2 4M (2070x2167)"
EIGER2_4M = Geometry("eiger2_4m", 2070, 2167, 1030, 514, 2, 4, 10, 37)
# BASELINE.json config 5
JUNGFRAU_4M = Geometry("jungfrau_4m", 2068, 2164, 1030, 514, 2, 4, 8, 36)

GEOMETRIES = {g.name: g for g in (EIGER2_16M, EIGER2_4M, JUNGFRAU_4M)}


def small_geometry(width: int, height: int, n_fast: int = 2, n_slow: int = 2, gap_fast: int = 5,
                   gap_slow: int = 7) -> Geometry:
    """A scaled-down detector for tests; modules need not tile the frame exactly
    (the remainder is added to the last module)."""
    mod_fast = (width - (n_fast - 1) * gap_fast) // n_fast
    mod_slow = (height - (n_slow - 1) * gap_slow) // n_slow
    return Geometry(f"small_{width}x{height}", width, height, mod_fast, mod_slow, n_fast, n_slow, gap_fast,
                    gap_slow)


def module_mask(g: Geometry) -> np.ndarray:
    """u8 mask, 1 on modules, 0 in the gaps (cf. h5read.c:1131-1156)."""
    m = np.ones((g.height, g.width), np.uint8)
    for i in range(1, g.n_slow):
        y = i * g.mod_slow + (i - 1) * g.gap_slow
        m[y:y + g.gap_slow, :] = 0
    for i in range(1, g.n_fast):
        x = i * g.mod_fast + (i - 1) * g.gap_fast
        m[:, x:x + g.gap_fast] = 0
    return m


def make_mask(g: Geometry, seed: int = BASE_SEED, dead_fraction: float = 1e-4) -> np.ndarray:
    rng = np.random.Generator(np.random.PCG64(seed))
    m = module_mask(g)
    n_dead = int(round(g.pixels * dead_fraction))
    if n_dead:
        idx = rng.integers(0, g.pixels, n_dead)
        m.reshape(-1)[idx] = 0
    return m


def make_frame(g: Geometry, mask: np.ndarray, dtype=np.uint16, config: int = 2, frame_index: int = 0,
               background: float = 2.0, n_spots: int | None = None, n_hot: int = 20,
               hot_value: int = 60000, spots: np.ndarray | None = None) -> np.ndarray:
    """One frame.  `spots` (optional) is an (N, 4) array of (x, y, amplitude, sigma)
    overriding the random spot list (used by sweeps)."""
    dtype = np.dtype(dtype)
    rng = np.random.Generator(np.random.PCG64(BASE_SEED + 1000 * config + frame_index))
    h, w = g.height, g.width
    img = rng.poisson(background, (h, w)).astype(np.int64)
    if spots is None:
        if n_spots is None:
            n_spots = max(1, int(round(300 * g.pixels / EIGER2_16M.pixels)))
        sx = rng.uniform(-1.0, w + 1.0, n_spots)
        sy = rng.uniform(-1.0, h + 1.0, n_spots)
        amp = np.exp(rng.uniform(np.log(20.0), np.log(5000.0), n_spots))
        sig = rng.uniform(0.8, 2.0, n_spots)
        spots = np.stack([sx, sy, amp, sig], axis=1)
    for x0, y0, a, s in spots:
        r = int(np.ceil(4 * s)) + 1
        xa, xb = max(0, int(x0) - r), min(w, int(x0) + r + 1)
        ya, yb = max(0, int(y0) - r), min(h, int(y0) + r + 1)
        if xa >= xb or ya >= yb:
            continue
        xs = np.arange(xa, xb) + 0.5 - x0
        ys = np.arange(ya, yb) + 0.5 - y0
        prof = a * np.exp(-0.5 * (ys[:, None] ** 2 + xs[None, :] ** 2) / (s * s))
        img[ya:yb, xa:xb] += rng.poisson(prof)
    if n_hot:
        hx = rng.integers(0, w, n_hot)
        hy = rng.integers(0, h, n_hot)
        img[hy, hx] = hot_value
    type_max = np.iinfo(dtype).max
    np.clip(img, 0, type_max, out=img)
    out = img.astype(dtype)
    # gaps and dead pixels read back as the type maximum on real detectors
    out[mask == 0] = type_max
    return out


def make_frames(g: Geometry, mask: np.ndarray, n: int, dtype=np.uint16, config: int = 2, **kw) -> np.ndarray:
    return np.stack([make_frame(g, mask, dtype, config, i, **kw) for i in range(n)])


def make_sweep(g: Geometry, mask: np.ndarray, n_frames: int, dtype=np.uint16, config: int = 4,
               n_reflections: int | None = None, background: float = 2.0) -> np.ndarray:
    """Rotation sweep: each reflection has a centre frame and a Gaussian rocking
    width, so spots span several consecutive frames (cfg 4)."""
    rng = np.random.Generator(np.random.PCG64(BASE_SEED + 1000 * config + 999_999))
    if n_reflections is None:
        n_reflections = max(4, int(round(40000 * (n_frames / 3600) * g.pixels / EIGER2_16M.pixels)))
    rx = rng.uniform(0, g.width, n_reflections)
    ry = rng.uniform(0, g.height, n_reflections)
    rz = rng.uniform(0, n_frames, n_reflections)
    amp = np.exp(rng.uniform(np.log(50.0), np.log(5000.0), n_reflections))
    sig = rng.uniform(0.8, 2.0, n_reflections)
    sz = rng.uniform(0.3, 1.5, n_reflections)
    frames = []
    for f in range(n_frames):
        zc = f + 0.5
        wgt = np.exp(-0.5 * ((zc - rz) / sz) ** 2)
        sel = wgt > 1e-3
        spots = np.stack([rx[sel], ry[sel], amp[sel] * wgt[sel], sig[sel]], axis=1)
        frames.append(make_frame(g, mask, dtype, config, f, background=background, n_hot=0, spots=spots))
    return np.stack(frames)


def domain_ok(frame: np.ndarray, mask: np.ndarray) -> bool:
    """Exactness domain of the double-precision SAT oracle:
    all valid pixels < 2^24 and sum of squares of valid pixels < 2^53."""
    v = frame[mask != 0].astype(np.float64)
    return bool((v < 2 ** 24).all() and float((v * v).sum()) < 2.0 ** 53)
