"""ctypes bindings for the CPU oracle (oracle/ffs_oracle.c) and, when built, the
reference's own standalone.cc (oracle/_ref/libffs_ref.so).

TEST INFRASTRUCTURE ONLY -- see oracle/ffs_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None
_REF = None

ALGO_DISPERSION = 0
ALGO_DISPERSION_EXTENDED = 1


class OracleParams(C.Structure):
    _fields_ = [
        ("kernel_half_x", C.c_int32),
        ("kernel_half_y", C.c_int32),
        ("min_count", C.c_int32),
        ("nsig_b", C.c_double),
        ("nsig_s", C.c_double),
        ("threshold", C.c_double),
    ]


class OracleSpot(C.Structure):
    _fields_ = [
        ("label", C.c_uint64),
        ("x_min", C.c_uint32), ("x_max", C.c_uint32),
        ("y_min", C.c_uint32), ("y_max", C.c_uint32),
        ("z_min", C.c_int32), ("z_max", C.c_int32),
        ("num_pixels", C.c_uint32),
        ("peak_x", C.c_uint32), ("peak_y", C.c_uint32),
        ("peak_z", C.c_int32),
        ("peak_intensity", C.c_uint64),
        ("sum_intensity", C.c_uint64),
        ("com_x", C.c_float), ("com_y", C.c_float), ("com_z", C.c_float),
        ("peak_centroid_distance", C.c_float),
    ]


SPOT_DTYPE = np.dtype(OracleSpot)


def build(force: bool = False) -> None:
    """Compile libffs_oracle.so (and _ref/libffs_ref.so where /root/reference exists)."""
    need = force or not (_HERE / "libffs_oracle.so").exists()
    src_newer = False
    so = _HERE / "libffs_oracle.so"
    if so.exists():
        src_newer = max((_HERE / f).stat().st_mtime for f in ("ffs_oracle.c", "ffs_oracle.h")) > so.stat().st_mtime
    if need or src_newer or (Path("/root/reference").exists() and not (_HERE / "_ref/libffs_ref.so").exists()):
        subprocess.run(["make", "-C", str(_HERE)], check=True, stdout=subprocess.DEVNULL)


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        build()
        L = C.CDLL(str(_HERE / "libffs_oracle.so"))
        P = C.POINTER(OracleParams)
        vp, u8p, i64p = C.c_void_p, C.c_void_p, C.c_void_p
        for name in ("ffs_oracle_dispersion", "ffs_oracle_dispersion_extended",
                     "ffs_oracle_dispersion_bruteforce", "ffs_oracle_dispersion_extended_bruteforce"):
            f = getattr(L, name)
            f.restype = C.c_int64
            f.argtypes = [vp, C.c_int, u8p, C.c_int, C.c_int, P, u8p]
        L.ffs_oracle_extended_stage.restype = C.c_int64
        L.ffs_oracle_extended_stage.argtypes = [vp, C.c_int, u8p, C.c_int, C.c_int, P, C.c_int, u8p]
        L.ffs_oracle_default_params.argtypes = [P]
        L.ffs_oracle_resolution_mask.restype = None
        L.ffs_oracle_resolution_mask.argtypes = [u8p, C.c_int, C.c_int] + [C.c_float] * 8
        L.ffs_oracle_connected_components.restype = C.c_int64
        L.ffs_oracle_connected_components.argtypes = [u8p, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                                      i64p, vp, C.c_int64]
        L.ffs_oracle_filter_spots.restype = C.c_int64
        L.ffs_oracle_filter_spots.argtypes = [vp, C.c_int64, C.c_uint32, C.c_float, i64p, i64p]
        L.ffs_oracle_process_frame.restype = C.c_int
        L.ffs_oracle_process_frame.argtypes = [vp, C.c_int, u8p, C.c_int, C.c_int, P, C.c_int, C.c_uint32,
                                               C.c_float, C.c_int, i64p]
        _LIB = L
    return _LIB


def ref_available() -> bool:
    return (_HERE / "_ref/libffs_ref.so").exists()


def ref() -> C.CDLL:
    """The reference's own standalone.cc behind oracle/ref_shim.cc."""
    global _REF
    if _REF is None:
        build()
        R = C.CDLL(str(_HERE / "_ref/libffs_ref.so"))
        R.ffs_ref_create.restype = C.c_void_p
        R.ffs_ref_create.argtypes = [C.c_size_t, C.c_size_t]
        R.ffs_ref_free.argtypes = [C.c_void_p]
        R.ffs_ref_standard_dispersion.restype = C.c_int64
        R.ffs_ref_standard_dispersion.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        R.ffs_ref_run_frames.restype = C.c_double
        R.ffs_ref_run_frames.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t,
                                         C.c_size_t, C.c_int, C.c_void_p]
        _REF = R
    return _REF


def default_params(**kw) -> OracleParams:
    p = OracleParams()
    lib().ffs_oracle_default_params(C.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def _prep(image: np.ndarray, mask: np.ndarray):
    assert image.ndim == 2 and image.dtype in (np.uint16, np.uint32), image.dtype
    image = np.ascontiguousarray(image)
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    assert mask.shape == image.shape
    h, w = image.shape
    return image, mask, w, h, image.dtype.itemsize


def _threshold(fn_name: str, image, mask, params, *extra):
    image, mask, w, h, bpp = _prep(image, mask)
    p = params or default_params()
    dst = np.empty((h, w), np.uint8)
    n = getattr(lib(), fn_name)(image.ctypes.data, bpp, mask.ctypes.data, w, h, C.byref(p), *extra, dst.ctypes.data)
    if n < 0:
        raise ValueError(f"{fn_name} failed: {n}")
    assert n == int(dst.sum())
    return dst


def dispersion(image, mask, params=None) -> np.ndarray:
    return _threshold("ffs_oracle_dispersion", image, mask, params)


def dispersion_extended(image, mask, params=None) -> np.ndarray:
    return _threshold("ffs_oracle_dispersion_extended", image, mask, params)


def extended_stage(image, mask, stage: int, params=None) -> np.ndarray:
    return _threshold("ffs_oracle_extended_stage", image, mask, params, stage)


def dispersion_bruteforce(image, mask, params=None) -> np.ndarray:
    return _threshold("ffs_oracle_dispersion_bruteforce", image, mask, params)


def dispersion_extended_bruteforce(image, mask, params=None) -> np.ndarray:
    return _threshold("ffs_oracle_dispersion_extended_bruteforce", image, mask, params)


def reference_dispersion(image, mask) -> np.ndarray:
    """Run the REFERENCE's standalone.cc (fixed parameters 3x3/2/0/6/3)."""
    image, mask, w, h, bpp = _prep(image, mask)
    R = ref()
    ctx = R.ffs_ref_create(w, h)
    try:
        dst = np.empty((h, w), np.uint8)
        n = R.ffs_ref_standard_dispersion(ctx, image.ctypes.data, bpp, mask.ctypes.data, dst.ctypes.data)
        assert n == int(dst.sum())
        return dst
    finally:
        R.ffs_ref_free(ctx)


def resolution_mask(mask, wavelength, distance_m, bcx_px, bcy_px, psx_m, psy_m, dmin, dmax) -> np.ndarray:
    out = np.ascontiguousarray(mask, dtype=np.uint8).copy()
    h, w = out.shape
    lib().ffs_oracle_resolution_mask(out.ctypes.data, w, h, wavelength, distance_m, bcx_px, bcy_px, psx_m, psy_m,
                                     dmin, dmax)
    return out


def connected_components(strong, image, row_wrap=True, want_labels=True):
    """strong/image: (H, W) for 2D or (Z, H, W) for 3D.  Returns (labels|None, spots)."""
    strong = np.ascontiguousarray(strong, dtype=np.uint8)
    image = np.ascontiguousarray(image)
    if strong.ndim == 2:
        strong3, image3 = strong[None], image[None]
    else:
        strong3, image3 = strong, image
    z, h, w = strong3.shape
    assert image3.shape == strong3.shape
    cap = max(int(np.count_nonzero(strong3)), 1)
    spots = np.zeros(cap, SPOT_DTYPE)
    labels = np.empty(strong3.shape, np.int64) if want_labels else None
    n = lib().ffs_oracle_connected_components(
        strong3.ctypes.data, image3.ctypes.data, image3.dtype.itemsize, w, h, z, int(bool(row_wrap)),
        labels.ctypes.data if want_labels else None, spots.ctypes.data, cap)
    if n < 0:
        raise ValueError(f"connected_components failed: {n}")
    if want_labels and strong.ndim == 2:
        labels = labels[0]
    return labels, spots[:n].copy()


def filter_spots(spots, min_spot_size=3, max_peak_centroid_separation=2.0):
    s = np.ascontiguousarray(spots).copy()
    rs, rp = C.c_int64(0), C.c_int64(0)
    kept = lib().ffs_oracle_filter_spots(s.ctypes.data, len(s), min_spot_size, max_peak_centroid_separation,
                                         C.addressof(rs), C.addressof(rp))
    return s[:kept].copy(), rs.value, rp.value


def process_frame(image, mask, params=None, algorithm=ALGO_DISPERSION, min_spot_size=3,
                  max_peak_centroid_separation=2.0, row_wrap=True):
    image, mask, w, h, bpp = _prep(image, mask)
    p = params or default_params()
    counts = (C.c_int64 * 4)()
    rc = lib().ffs_oracle_process_frame(image.ctypes.data, bpp, mask.ctypes.data, w, h, C.byref(p), algorithm,
                                        min_spot_size, max_peak_centroid_separation, int(bool(row_wrap)),
                                        C.addressof(counts))
    if rc:
        raise ValueError(f"process_frame failed: {rc}")
    return dict(num_strong_pixels=counts[0], n_spots_total=counts[1], num_strong_pixels_filtered=counts[2],
                n_spots_filtered=counts[3])
