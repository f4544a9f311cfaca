"""ctypes binding of libffs_b200.so (C ABI: include/ffs_b200.h).

The library is CUDA-only (sm_100a).  Loading works without a GPU (so the exported
symbols can be checked on a CPU box); every compute entry point fails loudly when there
is no B200: there is no CPU fallback anywhere in this package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_PKG = Path(__file__).resolve().parent
_ROOT = _PKG.parent
# FFS_B200_LIB: developer override to A/B a differently-built copy of the same library
LIB_PATH = Path(os.environ["FFS_B200_LIB"]) if os.environ.get("FFS_B200_LIB") else _PKG / "libffs_b200.so"

FFS_OK = 0
ALGO_DISPERSION = 0
ALGO_DISPERSION_EXTENDED = 1

STATUS = {0: "FFS_OK", -1: "FFS_ERR_ARG", -2: "FFS_ERR_CUDA", -3: "FFS_ERR_NO_DEVICE",
          -4: "FFS_ERR_CAPACITY", -5: "FFS_ERR_STATE", -6: "FFS_ERR_NOMEM", -7: "FFS_ERR_FORMAT"}


class FfsError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str = ""):
        self.status = status
        super().__init__(f"{where}: {STATUS.get(status, status)} {detail}".strip())


class Params(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32), ("bytes_per_pixel", C.c_uint32),
        ("algorithm", C.c_int32), ("kernel_half", C.c_int32), ("min_count", C.c_int32),
        ("nsig_b", C.c_double), ("nsig_s", C.c_double), ("threshold", C.c_double),
        ("min_spot_size", C.c_uint32), ("min_spot_size_3d", C.c_uint32),
        ("max_peak_centroid_separation", C.c_float),
        ("cc_row_wrap", C.c_int32), ("device", C.c_int32), ("n_slots", C.c_int32),
        ("max_strong_per_frame", C.c_uint32), ("max_spots_per_frame", C.c_uint32),
        ("trusted_max", C.c_int64),
    ]


class Spot(C.Structure):
    _fields_ = [
        ("label", C.c_uint64),
        ("x_min", C.c_uint32), ("x_max", C.c_uint32), ("y_min", C.c_uint32), ("y_max", C.c_uint32),
        ("z_min", C.c_int32), ("z_max", C.c_int32),
        ("num_pixels", C.c_uint32),
        ("peak_x", C.c_uint32), ("peak_y", C.c_uint32), ("peak_z", C.c_int32),
        ("peak_intensity", C.c_uint64), ("sum_intensity", C.c_uint64),
        ("com_x", C.c_float), ("com_y", C.c_float), ("com_z", C.c_float),
        ("peak_centroid_distance", C.c_float),
    ]


SPOT_DTYPE = np.dtype(Spot)


class KabschGeometry(C.Structure):
    """ffs_kabsch_geometry: what spotfinder.cc:1157-1168 builds its Panel / s0 / Scan from."""
    _fields_ = [
        ("distance_mm", C.c_double),
        ("beam_center_x_px", C.c_double), ("beam_center_y_px", C.c_double),
        ("pixel_size_x_mm", C.c_double), ("pixel_size_y_mm", C.c_double),
        ("wavelength", C.c_double),
        ("oscillation_width_deg", C.c_double),
    ]


class FrameResult(C.Structure):
    _fields_ = [
        ("frame_id", C.c_uint64),
        ("num_strong_pixels", C.c_uint32), ("n_components", C.c_uint32),
        ("n_spots_total", C.c_uint32), ("num_strong_pixels_filtered", C.c_uint32),
        ("n_spots_filtered", C.c_uint32), ("status", C.c_int32),
    ]


class ImageT(C.Structure):  # h5read.h:34-39
    _fields_ = [("data", C.c_void_p), ("mask", C.c_void_p), ("slow", C.c_size_t), ("fast", C.c_size_t)]


# every symbol include/ffs_b200.h declares: name -> (restype, argtypes)
_vp, _i, _sz = C.c_void_p, C.c_int, C.c_size_t
SYMBOLS = {
    "ffs_default_params": (None, [C.POINTER(Params)]),
    "ffs_create": (_i, [C.POINTER(Params), _vp, C.POINTER(_vp)]),
    "ffs_destroy": (None, [_vp]),
    "ffs_last_error": (C.c_char_p, [_vp]),
    "ffs_version": (C.c_char_p, []),
    "ffs_device_count": (_i, []),
    "ffs_device_description": (_i, [_i, C.c_char_p, _sz]),
    "ffs_apply_resolution_mask": (_i, [_vp] + [C.c_float] * 8),
    "ffs_get_mask": (_i, [_vp, _vp]),
    "ffs_mask_valid_count": (C.c_uint64, [_vp]),
    "ffs_upload": (_i, [_vp, _i, _i, _vp]),
    "ffs_threshold": (_i, [_vp, _i, _i, _vp]),
    "ffs_process": (_i, [_vp, _i, _i, _vp]),
    "ffs_sync": (_i, [_vp, _vp]),
    "ffs_get_strong_bits": (_i, [_vp, _i, _vp]),
    "ffs_get_strong_u8": (_i, [_vp, _i, _vp]),
    "ffs_get_strong_u8_static": (_i, [_vp, _i, _vp]),
    "ffs_noalloc_max_pixels": (_sz, []),
    "ffs_get_extended_stage_u8": (_i, [_vp, _i, _i, _vp]),
    "ffs_fetch": (_i, [_vp, _i, C.POINTER(FrameResult), _vp, _sz, C.POINTER(_sz)]),
    "ffs_get_labels": (_i, [_vp, _i, _vp]),
    "ffs_acquire": (_i, [_vp, C.POINTER(_i), C.POINTER(_vp)]),
    "ffs_submit": (_i, [_vp, _i, C.c_uint64]),
    "ffs_release": (_i, [_vp, _i]),
    "ffs_submit_copy": (_i, [_vp, C.c_uint64, _vp]),
    "ffs_wait": (_i, [_vp, C.POINTER(FrameResult), _vp, _sz, C.POINTER(_sz)]),
    "ffs_pending": (_i, [_vp]),
    "ffs_acquire_compressed": (_i, [_vp, C.POINTER(_i), C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_submit_compressed": (_i, [_vp, _i, C.c_uint64, _sz]),
    "ffs_slot_buffers": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_decode_chunk_device": (_i, [_vp, _i, _vp, _sz, _vp]),
    "ffs_sweep_begin": (_i, [_vp, C.c_uint64]),
    "ffs_sweep_record_slot": (_i, [_vp, _i, C.c_uint64]),
    "ffs_sweep_finish": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_sweep_export_plane": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_sweep_export_sums": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_comm_unique_id": (_i, [_vp]),
    "ffs_comm_create": (_i, [_vp, _i, _i, _i, C.POINTER(_vp)]),
    "ffs_comm_destroy": (None, [_vp]),
    "ffs_comm_rank": (_i, [_vp]),
    "ffs_comm_world": (_i, [_vp]),
    "ffs_comm_last_error": (C.c_char_p, [_vp]),
    "ffs_sweep_merge": (_i, [_vp, _vp, C.c_uint64, C.c_uint64, _vp, _sz, C.POINTER(_vp), C.POINTER(_sz),
                             C.POINTER(C.c_double)]),
    "ffs_sweep_gather": (_i, [_vp, _vp, _sz, C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_sweep_merge_host": (_i, [_i, C.c_uint32, C.c_uint32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                  C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_set_gain": (_i, [_vp, _vp]),
    "ffs_keep_pixel_lists": (_i, [_vp, _i]),
    "ffs_self_check": (_i, [_vp, _i]),
    "ffs_last_self_check": (_i, [_vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "ffs_frame_timing": (_i, [_vp, _i]),
    "ffs_last_frame_timing": (_i, [_vp, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "ffs_last_pixel_list": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_last_frame": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_sweep_export_moments": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "ffs_kabsch_variances": (_i, [_vp, C.POINTER(KabschGeometry), _vp, _vp, _vp, _sz, _vp, _vp, _vp]),
    "ffs_free": (None, [_vp]),
    "ffs_filter_spots": (_sz, [_vp, _sz, C.c_uint32, C.c_float, C.POINTER(_sz), C.POINTER(_sz)]),
    "ffs_kernel_launches": (C.c_uint64, [_vp]),
    "ffs_profile_enable": (_i, [_vp, _i]),
    "ffs_profile_read": (_i, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]),
    "ffs_frame_pitch": (_sz, [_vp]),
    "ffs_frame_bytes_device": (_sz, [_vp]),
    "ffs_bslz4_decode_chunk": (_i, [_vp, _sz, _vp, _sz, _sz]),
    "spotfinder_create": (_vp, [_sz, _sz]),
    "spotfinder_free": (None, [_vp]),
    "spotfinder_standard_dispersion": (C.c_uint32, [_vp, C.POINTER(ImageT), C.POINTER(_vp)]),
    "spotfinder_create32": (_vp, [_sz, _sz]),
    "spotfinder_standard_dispersion32": (C.c_uint32, [_vp, C.POINTER(ImageT), C.POINTER(_vp)]),
}

_LIB = None


def build(force: bool = False) -> Path:
    """Compile libffs_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    srcs = list((_ROOT / "csrc").glob("*")) + [_ROOT.parent / "include" / "ffs_b200.h", _ROOT / "Makefile"]
    stale = (not LIB_PATH.exists()) or max(s.stat().st_mtime for s in srcs) > LIB_PATH.stat().st_mtime
    if force or stale:
        import shutil
        if shutil.which("nvcc") is None and not Path("/usr/local/cuda/bin/nvcc").exists():
            if LIB_PATH.exists():
                return LIB_PATH
            raise RuntimeError("nvcc not found and libffs_b200.so is not built")
        subprocess.run(["make", "-C", str(_ROOT)], check=True, stdout=subprocess.DEVNULL)
    return LIB_PATH


def load() -> C.CDLL:
    """Load the CUDA library.  Raises if it is missing: nothing in this package computes
    on the CPU."""
    global _LIB
    if _LIB is None:
        if not LIB_PATH.exists():
            build()
        lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
    return _LIB
