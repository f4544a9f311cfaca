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
    if so.exists() and (_HERE / "bslz4_oracle.c").stat().st_mtime > so.stat().st_mtime:
        src_newer = True
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
        L.ffs_oracle_dispersion_gain.restype = C.c_int64
        L.ffs_oracle_dispersion_gain.argtypes = [vp, C.c_int, u8p, C.c_int, C.c_int, P, vp, u8p]
        L.ffs_oracle_dispersion_extended_gain.restype = C.c_int64
        L.ffs_oracle_dispersion_extended_gain.argtypes = [vp, C.c_int, u8p, C.c_int, C.c_int, P, vp, C.c_int, u8p]
        L.ffs_oracle_extended_stage.restype = C.c_int64
        L.ffs_oracle_extended_stage.argtypes = [vp, C.c_int, u8p, C.c_int, C.c_int, P, C.c_int, u8p]
        L.ffs_oracle_default_params.argtypes = [P]
        L.ffs_oracle_resolution_mask.restype = None
        L.ffs_oracle_resolution_mask.argtypes = [u8p, C.c_int, C.c_int] + [C.c_float] * 8
        L.ffs_oracle_bslz4_decode.restype = C.c_int
        L.ffs_oracle_bslz4_decode.argtypes = [vp, C.c_size_t, vp, C.c_size_t, C.c_size_t]
        L.ffs_oracle_chebyshev_to_zero.restype = None
        L.ffs_oracle_chebyshev_to_zero.argtypes = [u8p, C.c_int, C.c_int, vp]
        L.ffs_oracle_erode.restype = C.c_int
        L.ffs_oracle_erode.argtypes = [u8p, u8p, C.c_int, C.c_int, P]
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


_REFGPU = {}


def refgpu_available(bytes_per_pixel: int = 2) -> bool:
    return (_HERE / ("_ref/libffs_refgpu.so" if bytes_per_pixel == 2 else "_ref/libffs_refgpu32.so")).exists()


def refgpu(bytes_per_pixel: int = 2) -> C.CDLL:
    """The reference's own CUDA kernels (spotfinder/spotfinder.cu, kernels/*.cu) compiled in place
    for sm_100 behind oracle/refgpu_shim.cu: the "kernel to beat" and the resolution-mask pin."""
    if bytes_per_pixel not in _REFGPU:
        name = "_ref/libffs_refgpu.so" if bytes_per_pixel == 2 else "_ref/libffs_refgpu32.so"
        G = C.CDLL(str(_HERE / name))
        G.ffs_refgpu_last_error.restype = C.c_char_p
        G.ffs_refgpu_run.restype = C.c_int
        G.ffs_refgpu_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                     C.c_float, C.c_double, C.c_int, C.c_void_p, C.POINTER(C.c_double)]
        G.ffs_refgpu_run_kernels.restype = C.c_int
        G.ffs_refgpu_run_kernels.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_float, C.c_float, C.c_double, C.c_int, C.POINTER(C.c_double)]
        G.ffs_refgpu_resolution_mask.restype = C.c_int
        G.ffs_refgpu_resolution_mask.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_float] * 8
        assert G.ffs_refgpu_bytes_per_pixel() == bytes_per_pixel
        _REFGPU[bytes_per_pixel] = G
    return _REFGPU[bytes_per_pixel]


def refgpu_run(frames, mask, algorithm=ALGO_DISPERSION, min_count=3, nsig_b=6.0, nsig_s=3.0, max_valid=None,
               repeats=1, want_strong=True, kernels_only=False):
    """Run the reference GPU wrappers (call_do_spotfinding_dispersion/_extended,
    spotfinder/spotfinder.cuh:9-34; defaults min_count=3, 6, 3 as declared there) over
    frames (N, H, W).  Returns (strong (N,H,W) u8 | None, ms for `repeats` passes over all N)."""
    frames = np.ascontiguousarray(frames)
    if frames.ndim == 2:
        frames = frames[None]
    n, h, w = frames.shape
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    G = refgpu(frames.dtype.itemsize)
    if max_valid is None:
        max_valid = float(np.iinfo(frames.dtype).max)
    ms = C.c_double(0)
    if kernels_only:
        rc = G.ffs_refgpu_run_kernels(frames.ctypes.data, mask.ctypes.data, w, h, n, algorithm, min_count, nsig_b,
                                      nsig_s, max_valid, repeats, C.byref(ms))
        strong = None
    else:
        strong = np.empty((n, h, w), np.uint8) if want_strong else None
        rc = G.ffs_refgpu_run(frames.ctypes.data, mask.ctypes.data, w, h, n, algorithm, min_count, nsig_b, nsig_s,
                              max_valid, repeats, strong.ctypes.data if want_strong else None, C.byref(ms))
    if rc:
        raise RuntimeError(f"refgpu: {G.ffs_refgpu_last_error().decode()} ({rc})")
    return strong, ms.value


def refgpu_resolution_mask(mask, wavelength, distance_m, bcx_px, bcy_px, psx_m, psy_m, dmin, dmax) -> np.ndarray:
    """apply_resolution_filtering (spotfinder.cc:110-137) = the reference's apply_resolution_mask
    kernel itself (kernels/masking.cu:99-147), run on the GPU."""
    out = np.ascontiguousarray(mask, dtype=np.uint8).copy()
    h, w = out.shape
    G = refgpu(2)
    rc = G.ffs_refgpu_resolution_mask(out.ctypes.data, w, h, wavelength, distance_m, bcx_px, bcy_px, psx_m, psy_m,
                                      dmin, dmax)
    if rc:
        raise RuntimeError(f"refgpu: {G.ffs_refgpu_last_error().decode()} ({rc})")
    return out


_REFEXT = None


def refext_available() -> bool:
    return (_HERE / "_ref/libffs_refext.so").exists()


def refext() -> C.CDLL:
    global _REFEXT
    if _REFEXT is None:
        _REFEXT = C.CDLL(str(_HERE / "_ref/libffs_refext.so"))
        _REFEXT.ffs_refext_threshold.restype = C.c_long
        _REFEXT.ffs_refext_threshold.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                                 C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_int,
                                                 C.c_void_p]
        _REFEXT.ffs_refext_create.restype = C.c_void_p
        _REFEXT.ffs_refext_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_double,
                                              C.c_double, C.c_double, C.c_int]
        _REFEXT.ffs_refext_destroy.argtypes = [C.c_void_p]
        _REFEXT.ffs_refext_run.restype = C.c_long
        _REFEXT.ffs_refext_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    return _REFEXT


def refext_threshold(image, mask, algorithm=ALGO_DISPERSION_EXTENDED, params=None, gain=None, stage=3) -> np.ndarray:
    """The reference's own CPU DispersionThreshold / DispersionExtendedThreshold (baseline/spotfinder/
    baseline.cpp:259-310,730-805), compiled in place behind oracle/refext_shim.cc (stand-in scitbx /
    DIALS headers: oracle/stub_dials).  stage 1 / 2 stop the extended algorithm after its first pass /
    after the erosion.  Returns the (H, W) u8 plane."""
    G = refext()
    image, mask, w, h, bpp = _prep(image, mask)
    p = params or default_params()
    g = None if gain is None else np.ascontiguousarray(gain, dtype=np.float64)
    out = np.zeros((h, w), np.uint8)
    n = G.ffs_refext_threshold(algorithm, image.ctypes.data, bpp, mask.ctypes.data, w, h, p.kernel_half_x, p.kernel_half_y,
                                     p.nsig_b, p.nsig_s, p.threshold, p.min_count,
                                     None if g is None else g.ctypes.data, stage, out.ctypes.data)
    if n < 0:
        raise ValueError("the reference's argument checks (DIALS_ASSERT) failed")
    return out


class RefCCSpot(C.Structure):
    _fields_ = [("x_min", C.c_uint32), ("x_max", C.c_uint32), ("y_min", C.c_uint32), ("y_max", C.c_uint32),
                ("z_min", C.c_int32), ("z_max", C.c_int32), ("num_pixels", C.c_int32), ("depth", C.c_int32),
                ("com_x", C.c_float), ("com_y", C.c_float), ("com_z", C.c_float), ("pad", C.c_float),
                ("sigma_b_variance", C.c_double), ("sigma_m_variance", C.c_double)]


class RefCCBox(C.Structure):
    _fields_ = [("l", C.c_uint32), ("t", C.c_uint32), ("r", C.c_uint32), ("b", C.c_uint32), ("num_pixels", C.c_int32)]


REFCC_SPOT_DTYPE = np.dtype(RefCCSpot)
REFCC_BOX_DTYPE = np.dtype(RefCCBox)
_REFCC = {}


def refcc_available(bytes_per_pixel: int = 2) -> bool:
    return (_HERE / ("_ref/libffs_refcc.so" if bytes_per_pixel == 2 else "_ref/libffs_refcc32.so")).exists()


def refcc(bytes_per_pixel: int = 2) -> C.CDLL:
    """The reference's own host-side ConnectedComponents / Reflection3D / filter_reflections /
    variances_in_kabsch_space (spotfinder/connected_components/connected_components.cc, dx2 Panel
    and Scan) compiled in place behind oracle/refcc_shim.cc (stand-in Boost.Graph / Eigen:
    oracle/stub_cc)."""
    if bytes_per_pixel not in _REFCC:
        name = "_ref/libffs_refcc.so" if bytes_per_pixel == 2 else "_ref/libffs_refcc32.so"
        G = C.CDLL(str(_HERE / name))
        G.ffs_refcc_new.restype = C.c_void_p
        G.ffs_refcc_free.argtypes = [C.c_void_p]
        G.ffs_refcc_add_slice.restype = C.c_int
        G.ffs_refcc_add_slice.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                          C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        G.ffs_refcc_boxes.restype = C.c_int
        G.ffs_refcc_boxes.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        G.ffs_refcc_find_2d.restype = C.c_long
        G.ffs_refcc_find_2d.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_float, C.c_void_p, C.c_size_t,
                                        C.POINTER(C.c_uint32)]
        G.ffs_refcc_find_3d.restype = C.c_long
        G.ffs_refcc_find_3d.argtypes = [C.c_void_p, C.c_uint32, C.c_float, C.c_void_p, C.c_size_t]
        G.ffs_refcc_variances.restype = C.c_long
        G.ffs_refcc_variances.argtypes = [C.c_void_p] + [C.c_double] * 8 + [C.c_int, C.c_void_p, C.c_size_t]
        assert G.ffs_refcc_pixel_bytes() == bytes_per_pixel
        _REFCC[bytes_per_pixel] = G
    return _REFCC[bytes_per_pixel]


class RefCC:
    """One run of the reference's connected-components code over a frame or a sweep.

        with RefCC(frames, strong, min_spot_size=3) as r:      # (Z, H, W) or (H, W)
            r.num_strong, r.num_strong_filtered, r.boxes       # per slice (get_boxes())
            r.find_2d(slice, 3, 2.0)  -> (spots after both filters, components before)
            r.find_3d(3, 2.0)         -> spots after both filters, in the reference's order
            r.variances(geometry)     -> the same spots with sigma_b / sigma_m variances and depth
    """

    def __init__(self, frames, strong, min_spot_size=3):
        frames = np.ascontiguousarray(frames)
        strong = np.ascontiguousarray(strong, dtype=np.uint8)
        if frames.ndim == 2:
            frames, strong = frames[None], strong[None]
        assert frames.shape == strong.shape and frames.dtype in (np.uint16, np.uint32)
        self.G = refcc(frames.dtype.itemsize)
        self.h = self.G.ffs_refcc_new()
        self.n_images = len(frames)
        self.cap = max(int(np.count_nonzero(strong)), 1)
        self.num_strong, self.num_strong_filtered, self.boxes = [], [], []
        z, hgt, w = frames.shape
        for f, s in zip(frames, strong):
            a, b, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
            idx = self.G.ffs_refcc_add_slice(self.h, s.ctypes.data, f.ctypes.data, w, hgt, min_spot_size,
                                             C.byref(a), C.byref(b), C.byref(c))
            if idx < 0:
                raise RuntimeError("reference ConnectedComponents threw")
            boxes = np.zeros(max(c.value, 1), REFCC_BOX_DTYPE)
            self.G.ffs_refcc_boxes(self.h, idx, boxes.ctypes.data, len(boxes))
            self.num_strong.append(a.value)
            self.num_strong_filtered.append(b.value)
            self.boxes.append(boxes[:c.value].copy())

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.G.ffs_refcc_free(self.h)
        self.h = None

    def find_2d(self, slice_index, min_spot_size=3, max_sep=2.0):
        out = np.zeros(self.cap, REFCC_SPOT_DTYPE)
        before = C.c_uint32()
        n = self.G.ffs_refcc_find_2d(self.h, slice_index, min_spot_size, max_sep, out.ctypes.data, len(out), C.byref(before))
        if n < 0:
            raise RuntimeError("reference find_2d_components threw")
        return out[:n].copy(), before.value

    def find_3d(self, min_spot_size=3, max_sep=2.0):
        out = np.zeros(self.cap, REFCC_SPOT_DTYPE)
        n = self.G.ffs_refcc_find_3d(self.h, min_spot_size, max_sep, out.ctypes.data, len(out))
        if n < 0:
            raise RuntimeError("reference find_3d_components threw")
        self._last = out[:n].copy()
        return self._last.copy()

    def variances(self, geometry, oscillation_start=0.0):
        g = geometry
        out = self._last.copy()
        n = self.G.ffs_refcc_variances(self.h, g["distance_mm"], g["beam_center_x_px"], g["beam_center_y_px"],
                                       g["pixel_size_x_mm"], g["pixel_size_y_mm"], g["wavelength"], oscillation_start,
                                       g["oscillation_width_deg"], self.n_images, out.ctypes.data, len(out))
        if n < 0:
            raise RuntimeError("reference variances_in_kabsch_space threw")
        return out


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


def dispersion_gain(image, mask, gain, params=None) -> np.ndarray:
    """DispersionThreshold::threshold_w_gain (baseline.cpp:187-250,289-310); gain: (H, W) float64."""
    gain = np.ascontiguousarray(gain, dtype=np.float64)
    assert gain.shape == image.shape
    return _threshold("ffs_oracle_dispersion_gain", image, mask, params, gain.ctypes.data_as(C.c_void_p))


def dispersion_extended(image, mask, params=None) -> np.ndarray:
    return _threshold("ffs_oracle_dispersion_extended", image, mask, params)


def dispersion_extended_gain(image, mask, gain, params=None, stage: int = 3) -> np.ndarray:
    """DispersionExtendedThreshold::threshold_w_gain (baseline.cpp:484-541,657-714,772-805)."""
    gain = np.ascontiguousarray(gain, dtype=np.float64)
    assert gain.shape == image.shape
    return _threshold("ffs_oracle_dispersion_extended_gain", image, mask, params,
                      gain.ctypes.data_as(C.c_void_p), stage)


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


def bslz4_decode(chunk: bytes, dtype, n_elements: int) -> np.ndarray:
    """bshuf_decompress_lz4(chunk + 12, out, n, elem_size, 0) (spotfinder.cc:830-834), scalar C port."""
    out = np.empty(n_elements, dtype)
    rc = lib().ffs_oracle_bslz4_decode(chunk, len(chunk), out.ctypes.data, n_elements, out.dtype.itemsize)
    if rc:
        raise ValueError(f"bslz4_decode failed: {rc}")
    return out


def chebyshev_to_zero(plane) -> np.ndarray:
    """The oracle's restatement of dials chebyshev_distance(src, false, dst) (baseline.cpp:558)."""
    plane = np.ascontiguousarray(plane, dtype=np.uint8)
    h, w = plane.shape
    dist = np.empty((h, w), np.int32)
    lib().ffs_oracle_chebyshev_to_zero(plane.ctypes.data, w, h, dist.ctypes.data)
    return dist


def erode(mask, plane, params=None) -> np.ndarray:
    """erode_dispersion_mask (baseline.cpp:552-571): 1 = valid pixel outside the eroded signal core."""
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    out = np.ascontiguousarray(plane, dtype=np.uint8).copy()
    h, w = out.shape
    p = params or default_params()
    rc = lib().ffs_oracle_erode(mask.ctypes.data, out.ctypes.data, w, h, C.byref(p))
    if rc:
        raise ValueError(f"erode failed: {rc}")
    return out


def format_reflections_3d(spots) -> list[str]:
    """The lines of 3d_reflections.txt as the reference writes them (spotfinder.cc:1119-1148:
    ostream << with the default 6 significant digits for the float centre of mass; golden
    sample tests/3d_connected_components.sh:55-64)."""
    return [f"X: [{int(s['x_min'])}, {int(s['x_max'])}] Y: [{int(s['y_min'])}, {int(s['y_max'])}] "
            f"Z: [{int(s['z_min'])}, {int(s['z_max'])}] "
            f"COM: ({float(s['com_x']):g}, {float(s['com_y']):g}, {float(s['com_z']):g})" for s in spots]


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


def kabsch_variances(labels, image, spots, geometry):
    """Reflection3D::variances_in_kabsch_space (connected_components.cc:159-203) driven as the
    caller loop does (spotfinder.cc:1157-1196), restated pixel by pixel in fp64.  Pinned against
    the reference's own function (RefCC.variances, tests/test_oracle_refcc.py: 1e-12 relative).

    labels/image: (Z, H, W) from connected_components(); spots: its records (com_* as float32,
    the reference's center_of_mass() type).  geometry: dict with the ffs_kabsch_geometry keys.
    Returns (sigma_b_variance, sigma_m_variance, bbox_depth) arrays, one entry per spot.
    Geometry: dx2 Panel(distance, beam_center, pixel_size) with fast=+x, slow=-y
    (dx2/dx2/detector.cxx:243-268), px_to_mm without parallax (:193-198), get_lab_coord = d * (x, y, 1)
    (:211-213); s0 = (0, 0, -1/wavelength); rotation axis m2 = +x; Scan image range starts at 1."""
    g = geometry
    psx, psy = float(g["pixel_size_x_mm"]), float(g["pixel_size_y_mm"])
    fast = np.array([1.0, 0.0, 0.0])
    slow = np.array([0.0, -1.0, 0.0])
    origin = np.array([0.0, 0.0, -1.0 * float(g["distance_mm"])])
    origin = origin - float(g["beam_center_x_px"]) * psx * fast
    origin = origin - float(g["beam_center_y_px"]) * psy * slow
    s0 = np.array([0.0, 0.0, -1.0 / float(g["wavelength"])])
    m2 = np.array([1.0, 0.0, 0.0])
    osc_start, osc_width = 0.0, float(g["oscillation_width_deg"])
    image_range_0 = 1
    deg_to_rad = np.pi / 180.0

    def lab(xmm, ymm):
        return fast * xmm + slow * ymm + origin

    zz, yy, xx = np.nonzero(labels >= 0) if labels.dtype.kind == "i" else np.nonzero(labels)
    lab_of = labels[zz, yy, xx]
    order = np.argsort(lab_of, kind="stable")
    zz, yy, xx, lab_of = zz[order], yy[order], xx[order], lab_of[order]
    inten = image[zz, yy, xx].astype(np.float64)
    starts = np.searchsorted(lab_of, spots["label"].astype(lab_of.dtype), side="left")
    ends = np.searchsorted(lab_of, spots["label"].astype(lab_of.dtype), side="right")
    n = len(spots)
    sb, sm, depth = np.zeros(n), np.zeros(n), np.zeros(n, np.int32)
    for i, s in enumerate(spots):
        cx, cy, cz = float(s["com_x"]), float(s["com_y"]), float(s["com_z"])
        s1 = lab(cx * psx, cy * psy)
        phi = (osc_start + (cz - image_range_0) * osc_width) * deg_to_rad
        e1 = np.cross(s1, s0)
        e1 = e1 / np.sqrt(e1.dot(e1))
        e2 = np.cross(s1, e1)
        e2 = e2 / np.sqrt(e2.dot(e2))
        mags1 = np.sqrt(s1.dot(s1))
        zeta = m2.dot(e1)
        varx = vary = varz = tot = 0.0
        for k in range(starts[i], ends[i]):
            x, y, z = xx[k] + 0.5, yy[k] + 0.5, zz[k] + 0.5
            delta = lab(x * psx, y * psy) - s1
            eps1 = e1.dot(delta) / mags1
            eps2 = e2.dot(delta) / mags1
            phi_dash = (osc_start + (z - image_range_0) * osc_width) * deg_to_rad
            eps3 = (phi_dash - phi) * zeta
            varx += inten[k] * eps1 * eps1
            vary += inten[k] * eps2 * eps2
            varz += inten[k] * eps3 * eps3
            tot += inten[k]
        with np.errstate(invalid="ignore", divide="ignore"):
            varx, vary, varz = np.float64(varx) / tot, np.float64(vary) / tot, np.float64(varz) / tot
        sb[i] = (varx + vary) / 2.0
        sm[i] = varz
        depth[i] = int(s["z_max"]) - int(s["z_min"]) + 1
    return sb, sm, depth
