"""Host-side mirror of the reference's spot-finding interfaces over the C ABI.

Names follow the reference: `Spotfinder.standard_dispersion` is
StandaloneSpotfinder<T>::standard_dispersion (baseline/spotfinder/standalone.h:8-29),
`algorithm="dispersion" | "dispersion_extended"` is the CLI's -a/--algorithm
(spotfinder/spotfinder.cc:189-202,360-365), `ConnectedComponents`-style results
(boxes, num_strong_pixels, find_2d_components) come back from `process`/`fetch`.
All computation happens in libffs_b200.so on the GPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import lib as _lib
from .lib import (ALGO_DISPERSION, ALGO_DISPERSION_EXTENDED, FfsError, FrameResult, Params,
                  SPOT_DTYPE)

_ALGOS = {"dispersion": ALGO_DISPERSION, "dispersion_extended": ALGO_DISPERSION_EXTENDED}


def parse_algorithm(name: str) -> int:
    """Case-insensitive, as spotfinder.cc:189-202; unknown names raise ValueError
    (the reference throws std::invalid_argument)."""
    try:
        return _ALGOS[name.lower()]
    except KeyError:
        raise ValueError(f"Invalid algorithm specified: {name}") from None


def device_count() -> int:
    return _lib.load().ffs_device_count()


def device_description(device: int = 0) -> str:
    buf = C.create_string_buffer(256)
    rc = _lib.load().ffs_device_description(device, buf, 256)
    if rc:
        raise FfsError(rc, "ffs_device_description")
    return buf.value.decode()


class Spotfinder:
    """One context = one detector geometry + mask + parameter set on one GPU."""

    def __init__(self, width: int, height: int, dtype=np.uint16, mask: np.ndarray | None = None,
                 algorithm: str | int = "dispersion", n_slots: int = 4, device: int = 0, **params):
        self._L = _lib.load()
        self.dtype = np.dtype(dtype)
        if self.dtype not in (np.dtype(np.uint16), np.dtype(np.uint32)):
            raise ValueError("pixel type must be uint16 or uint32")
        p = Params()
        self._L.ffs_default_params(C.byref(p))
        p.width, p.height, p.bytes_per_pixel = width, height, self.dtype.itemsize
        p.algorithm = parse_algorithm(algorithm) if isinstance(algorithm, str) else int(algorithm)
        p.n_slots, p.device = n_slots, device
        for k, v in params.items():
            if not hasattr(p, k):
                raise AttributeError(f"unknown parameter {k}")
            setattr(p, k, v)
        self.params = p
        self.width, self.height, self.n_slots = width, height, n_slots
        m = None
        if mask is not None:
            m = np.ascontiguousarray(mask, dtype=np.uint8)
            if m.shape != (height, width):
                raise ValueError("mask shape must be (height, width)")
        ctx = C.c_void_p()
        rc = self._L.ffs_create(C.byref(p), m.ctypes.data if m is not None else None, C.byref(ctx))
        if rc:
            raise FfsError(rc, "ffs_create", (self._L.ffs_last_error(None) or b"").decode())
        self._ctx = ctx

    # -- plumbing --------------------------------------------------------------------
    def _ck(self, rc: int, where: str):
        if rc:
            raise FfsError(rc, where, (self._L.ffs_last_error(self._ctx) or b"").decode())

    def close(self):
        if getattr(self, "_ctx", None):
            self._L.ffs_destroy(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def kernel_launches(self) -> int:
        return int(self._L.ffs_kernel_launches(self._ctx))

    @property
    def mask_valid_count(self) -> int:
        return int(self._L.ffs_mask_valid_count(self._ctx))

    # -- mask ------------------------------------------------------------------------
    def apply_resolution_mask(self, wavelength, distance_m, beam_center_x_px, beam_center_y_px,
                              pixel_size_x_m, pixel_size_y_m, dmin=-1.0, dmax=-1.0):
        self._ck(self._L.ffs_apply_resolution_mask(self._ctx, wavelength, distance_m, beam_center_x_px,
                                                   beam_center_y_px, pixel_size_x_m, pixel_size_y_m,
                                                   dmin, dmax), "ffs_apply_resolution_mask")

    def get_mask(self) -> np.ndarray:
        out = np.empty((self.height, self.width), np.uint8)
        self._ck(self._L.ffs_get_mask(self._ctx, out.ctypes.data), "ffs_get_mask")
        return out

    # -- device-resident slot API ----------------------------------------------------------
    def upload(self, frames: np.ndarray, first_slot: int = 0) -> int:
        f = np.ascontiguousarray(frames, dtype=self.dtype)
        if f.ndim == 2:
            f = f[None]
        if f.shape[1:] != (self.height, self.width):
            raise ValueError("frame shape must be (height, width)")
        self._ck(self._L.ffs_upload(self._ctx, first_slot, f.shape[0], f.ctypes.data), "ffs_upload")
        return f.shape[0]

    def threshold(self, first_slot: int = 0, n: int = 1, stream: int | None = None, sync: bool = True):
        self._ck(self._L.ffs_threshold(self._ctx, first_slot, n, stream), "ffs_threshold")
        if sync:
            self.sync(stream)

    def process(self, first_slot: int = 0, n: int = 1, stream: int | None = None, sync: bool = True):
        self._ck(self._L.ffs_process(self._ctx, first_slot, n, stream), "ffs_process")
        if sync:
            self.sync(stream)

    def sync(self, stream: int | None = None):
        self._ck(self._L.ffs_sync(self._ctx, stream), "ffs_sync")

    def strong_bits(self, slot: int = 0) -> np.ndarray:
        out = np.empty((self.height, (self.width + 7) // 8), np.uint8)
        self._ck(self._L.ffs_get_strong_bits(self._ctx, slot, out.ctypes.data), "ffs_get_strong_bits")
        return out

    def strong_mask(self, slot: int = 0) -> np.ndarray:
        out = np.empty((self.height, self.width), np.uint8)
        self._ck(self._L.ffs_get_strong_u8(self._ctx, slot, out.ctypes.data), "ffs_get_strong_u8")
        return out

    def extended_stage(self, slot: int, stage: int) -> np.ndarray:
        out = np.empty((self.height, self.width), np.uint8)
        self._ck(self._L.ffs_get_extended_stage_u8(self._ctx, slot, stage, out.ctypes.data),
                 "ffs_get_extended_stage_u8")
        return out

    def fetch(self, slot: int = 0, max_spots: int | None = None):
        """(FrameResult, spots ndarray[SPOT_DTYPE]) of the last process() of this slot."""
        res = FrameResult()
        n = C.c_size_t(0)
        self._ck(self._L.ffs_fetch(self._ctx, slot, C.byref(res), None, 0, C.byref(n)), "ffs_fetch")
        cap = n.value if max_spots is None else min(max_spots, n.value)
        spots = np.zeros(max(cap, 1), SPOT_DTYPE)
        self._ck(self._L.ffs_fetch(self._ctx, slot, C.byref(res), spots.ctypes.data, cap, C.byref(n)),
                 "ffs_fetch")
        return res, spots[:cap]

    def labels(self, slot: int = 0) -> np.ndarray:
        out = np.empty((self.height, self.width), np.int64)
        self._ck(self._L.ffs_get_labels(self._ctx, slot, out.ctypes.data), "ffs_get_labels")
        return out

    # -- reference-shaped conveniences ---------------------------------------------------
    def standard_dispersion(self, image: np.ndarray) -> np.ndarray:
        """StandaloneSpotfinder::standard_dispersion: image -> bool (H, W) strong mask."""
        self.upload(image, 0)
        self.threshold(0, 1)
        return self.strong_mask(0).astype(bool)

    def filter_spots(self, spots: np.ndarray, min_spot_size: int | None = None,
                     max_peak_centroid_separation: float | None = None):
        """filter_reflections (connected_components.cc:207-236).  Returns
        (kept, n_removed_size, n_removed_separation)."""
        s = np.ascontiguousarray(spots).copy()
        a, b = C.c_size_t(0), C.c_size_t(0)
        ms = self.params.min_spot_size if min_spot_size is None else min_spot_size
        sep = self.params.max_peak_centroid_separation if max_peak_centroid_separation is None \
            else max_peak_centroid_separation
        kept = self._L.ffs_filter_spots(s.ctypes.data, len(s), ms, sep, C.byref(a), C.byref(b))
        return s[:kept].copy(), a.value, b.value

    # -- streaming API (host frames in, per-frame results out) -----------------------------
    def acquire(self):
        """(slot, ndarray view of the slot's pinned host frame) -- fill it, then submit()."""
        slot, ptr = C.c_int(-1), C.c_void_p()
        self._ck(self._L.ffs_acquire(self._ctx, C.byref(slot), C.byref(ptr)), "ffs_acquire")
        n = self.height * self.width
        buf = (C.c_uint8 * (n * self.dtype.itemsize)).from_address(ptr.value)
        return slot.value, np.frombuffer(buf, dtype=self.dtype).reshape(self.height, self.width)

    def submit(self, slot: int, frame_id: int):
        self._ck(self._L.ffs_submit(self._ctx, slot, frame_id), "ffs_submit")

    def submit_copy(self, frame_id: int, frame: np.ndarray):
        f = np.ascontiguousarray(frame, dtype=self.dtype)
        if f.shape != (self.height, self.width):
            raise ValueError("frame shape must be (height, width)")
        self._ck(self._L.ffs_submit_copy(self._ctx, frame_id, f.ctypes.data), "ffs_submit_copy")

    # -- compressed ingestion: bitshuffle-LZ4 chunk decoded on the device ------------------------
    def acquire_compressed(self):
        """(slot, uint8 view of the slot's pinned chunk buffer) -- fill it with the raw HDF5
        filter chunk (12-byte header + body), then submit_compressed(slot, id, n_bytes)."""
        slot, ptr, cap = C.c_int(-1), C.c_void_p(), C.c_size_t(0)
        self._ck(self._L.ffs_acquire_compressed(self._ctx, C.byref(slot), C.byref(ptr), C.byref(cap)),
                 "ffs_acquire_compressed")
        buf = (C.c_uint8 * cap.value).from_address(ptr.value)
        return slot.value, np.frombuffer(buf, dtype=np.uint8)

    def submit_compressed(self, slot: int, frame_id: int, n_bytes: int):
        self._ck(self._L.ffs_submit_compressed(self._ctx, slot, frame_id, n_bytes), "ffs_submit_compressed")

    def decode_chunk(self, chunk: bytes, slot: int = 0) -> np.ndarray:
        """Parity hook: decode one chunk on the device and return the dense frame."""
        out = np.empty((self.height, self.width), self.dtype)
        self._ck(self._L.ffs_decode_chunk_device(self._ctx, slot, chunk, len(chunk), out.ctypes.data),
                 "ffs_decode_chunk_device")
        return out

    @property
    def pending(self) -> int:
        return self._L.ffs_pending(self._ctx)

    def wait(self, max_spots: int = 4096):
        """Oldest submitted frame: (FrameResult, spots[:min(n, max_spots)])."""
        res, n = FrameResult(), C.c_size_t(0)
        spots = np.zeros(max(max_spots, 1), SPOT_DTYPE)
        self._ck(self._L.ffs_wait(self._ctx, C.byref(res), spots.ctypes.data, max_spots, C.byref(n)), "ffs_wait")
        return res, spots[:min(n.value, max_spots)]

    def set_gain(self, gain):
        """Per-pixel gain map (H, W) float64 for threshold_w_gain: standard dispersion
        (baseline.cpp:289-310) or dispersion_extended (baseline.cpp:772-805); None removes it."""
        if gain is None:
            self._ck(self._L.ffs_set_gain(self._ctx, None), "ffs_set_gain")
            return
        g = np.ascontiguousarray(gain, dtype=np.float64)
        if g.shape != (self.height, self.width):
            raise ValueError("gain must be (height, width)")
        self._ck(self._L.ffs_set_gain(self._ctx, g.ctypes.data), "ffs_set_gain")

    # -- rotation sweeps: 3D connected components --------------------------------------------
    def sweep_begin(self, max_frames: int = 0):
        self._ck(self._L.ffs_sweep_begin(self._ctx, max_frames), "ffs_sweep_begin")

    def sweep_record_slot(self, slot: int, frame_id: int):
        self._ck(self._L.ffs_sweep_record_slot(self._ctx, slot, frame_id), "ffs_sweep_record_slot")

    def sweep_finish(self) -> np.ndarray:
        """All 3D components of the recorded frames, unfiltered, canonical order
        (find_3d_components before filter_reflections)."""
        ptr, n = C.c_void_p(), C.c_size_t(0)
        self._ck(self._L.ffs_sweep_finish(self._ctx, C.byref(ptr), C.byref(n)), "ffs_sweep_finish")
        if not n.value:
            return np.zeros(0, SPOT_DTYPE)
        buf = (C.c_uint8 * (n.value * SPOT_DTYPE.itemsize)).from_address(ptr.value)
        out = np.frombuffer(buf, dtype=SPOT_DTYPE).copy()
        self._L.ffs_free(ptr)
        return out

    def sweep_export_plane(self, which: int) -> np.ndarray:
        """(n, 2) uint32 {linear index, component rank} of the first (0) / last (1) slice."""
        ptr, n = C.c_void_p(), C.c_size_t(0)
        self._ck(self._L.ffs_sweep_export_plane(self._ctx, which, C.byref(ptr), C.byref(n)),
                 "ffs_sweep_export_plane")
        if not n.value:
            return np.zeros((0, 2), np.uint32)
        buf = (C.c_uint32 * (2 * n.value)).from_address(ptr.value)
        out = np.frombuffer(buf, dtype=np.uint32).reshape(-1, 2).copy()
        self._L.ffs_free(ptr)
        return out

    # -- instrumentation -----------------------------------------------------------------------
    def profile_enable(self, on: bool = True):
        self._ck(self._L.ffs_profile_enable(self._ctx, int(on)), "ffs_profile_enable")

    def profile_read(self):
        """(total ms, launches) of the bracketed dominant kernel since the last read."""
        ms, n = C.c_double(0), C.c_uint64(0)
        self._ck(self._L.ffs_profile_read(self._ctx, C.byref(ms), C.byref(n)), "ffs_profile_read")
        return ms.value, n.value

    def sweep_export_sums(self) -> np.ndarray:
        """(n, 4) uint64 exact sums {I, (2x+1)I, (2y+1)I, (2z+1)I} of the last sweep_finish."""
        ptr, n = C.c_void_p(), C.c_size_t(0)
        self._ck(self._L.ffs_sweep_export_sums(self._ctx, C.byref(ptr), C.byref(n)), "ffs_sweep_export_sums")
        if not n.value:
            return np.zeros((0, 4), np.uint64)
        buf = (C.c_uint64 * (4 * n.value)).from_address(ptr.value)
        out = np.frombuffer(buf, dtype=np.uint64).reshape(-1, 4).copy()
        self._L.ffs_free(ptr)
        return out

    def sweep_export_moments(self) -> np.ndarray:
        """(n, 4) uint64 bbox-relative second moments {I u^2, I v^2, I u v, I w^2} of the last
        sweep_finish (u = x - x_min, v = y - y_min, w = z - z_min)."""
        ptr, n = C.c_void_p(), C.c_size_t(0)
        self._ck(self._L.ffs_sweep_export_moments(self._ctx, C.byref(ptr), C.byref(n)), "ffs_sweep_export_moments")
        if not n.value:
            return np.zeros((0, 4), np.uint64)
        buf = (C.c_uint64 * (4 * n.value)).from_address(ptr.value)
        out = np.frombuffer(buf, dtype=np.uint64).reshape(-1, 4).copy()
        self._L.ffs_free(ptr)
        return out

    def kabsch_variances(self, geometry: dict, spots: np.ndarray, sums: np.ndarray, moments: np.ndarray):
        """(sigma_b_variance, sigma_m_variance, bbox_depth) per spot -- the tuple
        Reflection3D::variances_in_kabsch_space returns (connected_components.cc:159-203),
        evaluated on the device.  geometry keys: distance_mm, beam_center_x_px, beam_center_y_px,
        pixel_size_x_mm, pixel_size_y_mm, wavelength, oscillation_width_deg."""
        from .lib import KabschGeometry
        g = KabschGeometry(**{k: float(v) for k, v in geometry.items()})
        spots = np.ascontiguousarray(spots, dtype=SPOT_DTYPE)
        sums = np.ascontiguousarray(sums, dtype=np.uint64).reshape(-1, 4)
        moments = np.ascontiguousarray(moments, dtype=np.uint64).reshape(-1, 4)
        n = len(spots)
        if len(sums) != n or len(moments) != n:
            raise ValueError("spots, sums and moments must describe the same spots")
        sb, sm, depth = np.zeros(n, np.float64), np.zeros(n, np.float64), np.zeros(n, np.int32)
        self._ck(self._L.ffs_kabsch_variances(self._ctx, C.byref(g), spots.ctypes.data, sums.ctypes.data,
                                              moments.ctypes.data, n, sb.ctypes.data, sm.ctypes.data,
                                              depth.ctypes.data), "ffs_kabsch_variances")
        return sb, sm, depth
