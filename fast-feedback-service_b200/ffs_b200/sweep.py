"""Slab-sharded rotation sweeps: merge per-GPU 3D components across slab boundaries.

The reference labels a whole sweep on one host after the last frame
(ConnectedComponents::find_3d_components, connected_components.cc:270-470).  Here each rank
owns a contiguous slab of frames and labels it on its GPU (ffs_sweep_finish); components can
only continue across a slab boundary through pixels with the SAME linear index in the last
frame of slab r and the first frame of slab r+1 (cc:352-370).  So the exchange is: every rank
publishes {its unfiltered spots, their exact integer sums, its first / last strong plane as
(linear index, component) pairs}; boundary planes are intersected, the touching components
are unioned, and merged spots are recomputed from the exact sums with the reference's own
arithmetic (double divide -> float centre of mass, float peak-centroid distance).  Canonical
label / output order = minimum (z, linear index), as on a single GPU.

The product's merge is native (csrc/api_multi.inl): `merge_slabs_native` = ffs_sweep_merge_host
(slabs in one address space), `NcclComm` + `merge_native` = ffs_sweep_merge (one rank per GPU,
boundary planes by ncclSend/ncclRecv, boundary-component records by ncclAllGather).
`merge_slabs` below is the numpy mirror kept as the test reference of that code (unit-tested on
CPU against the oracle's whole-sweep labelling); `gather_and_merge` moves whole slab tables with
torch.distributed (gloo in the CPU tests) and is no longer on the bench path.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .lib import SPOT_DTYPE


@dataclass
class SlabResult:
    z0: int                 # global index of the slab's first frame
    n_frames: int
    spots: np.ndarray       # SPOT_DTYPE, z local to the slab, canonical order
    sums: np.ndarray        # (n, 4) uint64: sum I, (2x+1)I, (2y+1)I, (2z+1)I  (z local)
    first: np.ndarray       # (k, 2) uint32 {linear index, component} of the slab's first frame
    last: np.ndarray        # (k, 2) uint32 of its last frame
    moments: np.ndarray | None = None  # (n, 4) uint64: sum I u^2, I v^2, I u v, I w^2 about the box minimum


def slab_bounds(n_frames: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous slabs: rank r owns frames [r*F/G, (r+1)*F/G)."""
    return rank * n_frames // world, (rank + 1) * n_frames // world


def _find(parent, i):
    while parent[i] != i:
        parent[i] = parent[parent[i]]
        i = parent[i]
    return i


def _absolute_moments(spots, sums, moments):
    """bbox-relative second moments -> moments about the origin (exact Python integers).
    sum I x^2 = sum I u^2 + 2 a sum I u + a^2 sum I, with sum I u = sum I x - a sum I, a = x_min."""
    out = np.zeros((len(spots), 4), dtype=object)
    for i in range(len(spots)):
        S0 = int(sums[i, 0])
        Ix, Iy, Iz = ((int(sums[i, k]) - S0) // 2 for k in (1, 2, 3))
        a, b, c = int(spots["x_min"][i]), int(spots["y_min"][i]), int(spots["z_min"][i])
        Iu, Iv, Iw = Ix - a * S0, Iy - b * S0, Iz - c * S0
        m = [int(v) for v in moments[i]]
        out[i, 0] = m[0] + 2 * a * Iu + a * a * S0
        out[i, 1] = m[1] + 2 * b * Iv + b * b * S0
        out[i, 2] = m[2] + a * Iv + b * Iu + a * b * S0
        out[i, 3] = m[3] + 2 * c * Iw + c * c * S0
    return out


def _relative_moments(spots, sums, absm):
    """Inverse of _absolute_moments for the (merged) bounding boxes in `spots`."""
    out = np.zeros((len(spots), 4), dtype=np.uint64)
    for i in range(len(spots)):
        S0 = int(sums[i, 0])
        Ix, Iy, Iz = ((int(sums[i, k]) - S0) // 2 for k in (1, 2, 3))
        a, b, c = int(spots["x_min"][i]), int(spots["y_min"][i]), int(spots["z_min"][i])
        out[i, 0] = int(absm[i, 0]) - 2 * a * Ix + a * a * S0
        out[i, 1] = int(absm[i, 1]) - 2 * b * Iy + b * b * S0
        out[i, 2] = int(absm[i, 2]) - a * Iy - b * Ix + a * b * S0
        out[i, 3] = int(absm[i, 3]) - 2 * c * Iz + c * c * S0
    return out


def merge_slabs(slabs: list[SlabResult], width: int, height: int, with_sums: bool = False):
    """All 3D components of the sweep (unfiltered, canonical order) from per-slab results.
    with_sums: also return the merged exact sums (n, 4) and, if every slab carries them, the
    merged bbox-relative second moments (n, 4) -- the inputs of Spotfinder.kabsch_variances."""
    slabs = sorted(slabs, key=lambda s: s.z0)
    px = width * height
    offs = np.cumsum([0] + [len(s.spots) for s in slabs])
    n = int(offs[-1])
    if n == 0:
        empty = np.zeros(0, SPOT_DTYPE)
        return (empty, np.zeros((0, 4), np.uint64), np.zeros((0, 4), np.uint64)) if with_sums else empty
    spots = np.concatenate([s.spots for s in slabs]).copy()
    sums = np.concatenate([s.sums.reshape(-1, 4) for s in slabs]).astype(object)  # exact big ints
    have_moments = all(s.moments is not None for s in slabs)
    absm = None
    if have_moments:  # about the origin with slab-local z, so that they simply add
        absm = np.concatenate([_absolute_moments(s.spots, s.sums.reshape(-1, 4), s.moments.reshape(-1, 4))
                               for s in slabs])
    # globalise z (slab-local -> sweep): bbox, peak, label, and sum (2z+1)I += 2*z0*sum I
    for s, a, b in zip(slabs, offs[:-1], offs[1:]):
        spots["z_min"][a:b] += s.z0
        spots["z_max"][a:b] += s.z0
        spots["peak_z"][a:b] += s.z0
        spots["label"][a:b] += np.uint64(s.z0 * px)
        if absm is not None:  # sum I (z + z0)^2, with sum I z = (sum (2z+1) I - sum I) / 2 still slab-local
            absm[a:b, 3] += 2 * s.z0 * ((sums[a:b, 3] - sums[a:b, 0]) // 2) + s.z0 * s.z0 * sums[a:b, 0]
        sums[a:b, 3] += 2 * s.z0 * sums[a:b, 0]
    parent = list(range(n))
    for k in range(len(slabs) - 1):
        lo, hi = slabs[k], slabs[k + 1]
        if lo.z0 + lo.n_frames != hi.z0 or not len(lo.last) or not len(hi.first):
            continue  # not adjacent frames, or nothing strong on one side
        _, ia, ib = np.intersect1d(lo.last[:, 0], hi.first[:, 0], assume_unique=True, return_indices=True)
        for ca, cb in set(zip((lo.last[ia, 1] + offs[k]).tolist(), (hi.first[ib, 1] + offs[k + 1]).tolist())):
            ra, rb = _find(parent, ca), _find(parent, cb)
            if ra != rb:  # smaller global id (= smaller (z, index)) stays root
                parent[max(ra, rb)] = min(ra, rb)
    roots = np.array([_find(parent, i) for i in range(n)])
    touched = np.flatnonzero(roots != np.arange(n))
    for i in touched:  # fold member i into its root r (r < i, so r keeps the canonical label)
        r = roots[i]
        a, b = spots[r], spots[i]
        for f in ("x_min", "y_min", "z_min"):
            a[f] = min(a[f], b[f])
        for f in ("x_max", "y_max", "z_max"):
            a[f] = max(a[f], b[f])
        a["num_pixels"] += b["num_pixels"]
        a["sum_intensity"] += b["sum_intensity"]
        # peak: brightest, ties -> smallest (z, y, x)  (connected_components.cc:143-157)
        ka = (-int(a["peak_intensity"]), int(a["peak_z"]), int(a["peak_y"]), int(a["peak_x"]))
        kb = (-int(b["peak_intensity"]), int(b["peak_z"]), int(b["peak_y"]), int(b["peak_x"]))
        if kb < ka:
            for f in ("peak_intensity", "peak_x", "peak_y", "peak_z"):
                a[f] = b[f]
        sums[r] += sums[i]
        if absm is not None:
            absm[r] += absm[i]
    # Reflection3D::center_of_mass / peak_centroid_distance, recomputed from the exact sums for
    # every surviving spot: merged ones changed, and slab-local z became sweep z for the rest
    # (vectorised; int -> float64 rounds to nearest exactly as the C++ static_cast<double> does)
    alive = np.flatnonzero(roots == np.arange(n))
    f64 = np.array([[float(v) for v in sums[r]] for r in alive], dtype=np.float64).reshape(-1, 4)
    with np.errstate(invalid="ignore", divide="ignore"):
        com = [((f64[:, k] * 0.5) / f64[:, 0]).astype(np.float32) for k in (1, 2, 3)]
        d = [(spots[f][alive].astype(np.float32) + np.float32(0.5)) - c
             for f, c in zip(("peak_x", "peak_y", "peak_z"), com)]
        dist = np.sqrt((d[0] * d[0] + d[1] * d[1]).astype(np.float32) + d[2] * d[2]).astype(np.float32)
    zero = f64[:, 0] == 0
    for arr in (*com, dist):
        arr[zero] = np.nan
    spots["com_x"][alive], spots["com_y"][alive], spots["com_z"][alive] = com
    spots["peak_centroid_distance"][alive] = dist
    keep = roots == np.arange(n)
    out = spots[keep]
    order = np.argsort(out["label"], kind="stable")
    if not with_sums:
        return out[order]
    out = out[order]
    sums_out = sums[keep][order]
    mom_out = _relative_moments(out, sums_out, absm[keep][order]) if absm is not None else None
    return out, sums_out.astype(np.uint64), mom_out


def gather_and_merge(local: SlabResult, width: int, height: int, group=None) -> np.ndarray:
    """Every rank contributes its slab; every rank returns the merged sweep."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return merge_slabs([local], width, height)
    gathered = [None] * dist.get_world_size(group)
    dist.all_gather_object(gathered, local, group=group)
    return merge_slabs(gathered, width, height)


def slab_result_from(sf, z0: int, n_frames: int) -> SlabResult:
    """Finish the sweep recorded in Spotfinder `sf` and package it for merging."""
    spots = sf.sweep_finish()
    return SlabResult(z0, n_frames, spots, sf.sweep_export_sums(), sf.sweep_export_plane(0),
                      sf.sweep_export_plane(1), sf.sweep_export_moments())


# ---- native merge (csrc/api_multi.inl) ----------------------------------------------------------
def merge_slabs_native(slabs: list[SlabResult], width: int, height: int, with_sums: bool = False):
    """ffs_sweep_merge_host: every slab's tables in this process -> the whole sweep's components."""
    import ctypes as C
    from . import lib as _lib
    L = _lib.load()
    n = len(slabs)
    keep = []  # keep the arrays alive across the call

    def arr(a, dtype):
        a = np.ascontiguousarray(a, dtype=dtype)
        keep.append(a)
        return a.ctypes.data

    z0 = arr([s.z0 for s in slabs], np.uint64)
    nf = arr([s.n_frames for s in slabs], np.uint64)
    have_m = all(s.moments is not None for s in slabs)
    P = C.c_void_p * n
    spots = P(*[arr(s.spots, SPOT_DTYPE) for s in slabs])
    sums = P(*[arr(np.asarray(s.sums).reshape(-1), np.uint64) for s in slabs])
    moms = P(*[arr(np.asarray(s.moments).reshape(-1), np.uint64) for s in slabs]) if have_m else None
    first = P(*[arr(np.asarray(s.first).reshape(-1), np.uint32) for s in slabs])
    last = P(*[arr(np.asarray(s.last).reshape(-1), np.uint32) for s in slabs])
    S = C.c_size_t * n
    n_spots = S(*[len(s.spots) for s in slabs])
    n_first = S(*[len(s.first) for s in slabs])
    n_last = S(*[len(s.last) for s in slabs])
    out, osum, omom, n_out = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_size_t(0)
    rc = L.ffs_sweep_merge_host(n, width, height, z0, nf, C.addressof(spots), C.addressof(n_spots), C.addressof(sums),
                                C.addressof(moms) if moms is not None else None, C.addressof(first),
                                C.addressof(n_first), C.addressof(last), C.addressof(n_last), C.byref(out),
                                C.byref(osum), C.byref(omom), C.byref(n_out))
    if rc:
        raise _lib.FfsError(rc, "ffs_sweep_merge_host")

    def take(ptr, dtype, count):
        if not ptr.value or not count:
            return np.zeros(count, dtype)
        a = np.frombuffer((C.c_uint8 * (count * np.dtype(dtype).itemsize)).from_address(ptr.value), dtype=dtype).copy()
        L.ffs_free(ptr)
        return a

    k = n_out.value
    res = take(out, SPOT_DTYPE, k)
    ssum = take(osum, np.uint64, 4 * k).reshape(-1, 4)
    smom = take(omom, np.uint64, 4 * k).reshape(-1, 4)
    return (res, ssum, smom if have_m else None) if with_sums else res


class NcclComm:
    """ffs_comm: one NCCL rank per GPU.  `NcclComm.from_torch()` bootstraps the unique id through
    the process group that torchrun already set up (rank 0 creates it, a broadcast delivers it)."""

    def __init__(self, unique_id: bytes, rank: int, world: int, device: int):
        import ctypes as C
        from . import lib as _lib
        self._L = _lib.load()
        h = C.c_void_p()
        rc = self._L.ffs_comm_create(unique_id, rank, world, device, C.byref(h))
        if rc:
            raise _lib.FfsError(rc, "ffs_comm_create", (self._L.ffs_comm_last_error(None) or b"").decode())
        self._h, self.rank, self.world = h, rank, world

    @staticmethod
    def unique_id() -> bytes:
        import ctypes as C
        from . import lib as _lib
        buf = (C.c_uint8 * 128)()
        rc = _lib.load().ffs_comm_unique_id(buf)
        if rc:
            raise _lib.FfsError(rc, "ffs_comm_unique_id")
        return bytes(buf)

    @classmethod
    def from_torch(cls, device: int):
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        t = torch.zeros(128, dtype=torch.uint8, device=f"cuda:{device}")
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(cls.unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        return cls(bytes(t.cpu().numpy().tobytes()), rank, world, device)

    def close(self):
        if getattr(self, "_h", None):
            self._L.ffs_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def merge_native(sf, comm: NcclComm, z0: int, n_frames: int, spots: np.ndarray, gather: bool = True):
    """ffs_sweep_merge (+ ffs_sweep_gather): `spots` = sf.sweep_finish() of this rank's slab.
    Returns (this rank's share | the whole sweep's list on every rank, exchange seconds)."""
    import ctypes as C
    from . import lib as _lib
    L = sf._L
    s = np.ascontiguousarray(spots, dtype=SPOT_DTYPE)
    out, n_out, secs = C.c_void_p(), C.c_size_t(0), C.c_double(0)
    rc = L.ffs_sweep_merge(sf._ctx, comm._h, z0, n_frames, s.ctypes.data, len(s), C.byref(out), C.byref(n_out),
                           C.byref(secs))
    if rc:
        raise _lib.FfsError(rc, "ffs_sweep_merge", (L.ffs_last_error(sf._ctx) or b"").decode())

    def take(ptr, count):
        if not ptr.value or not count:
            return np.zeros(0, SPOT_DTYPE)
        a = np.frombuffer((C.c_uint8 * (count * SPOT_DTYPE.itemsize)).from_address(ptr.value), dtype=SPOT_DTYPE).copy()
        L.ffs_free(ptr)
        return a

    mine = take(out, n_out.value)
    if not gather:
        return mine, secs.value
    allp, n_all = C.c_void_p(), C.c_size_t(0)
    rc = L.ffs_sweep_gather(comm._h, mine.ctypes.data if len(mine) else None, len(mine), C.byref(allp), C.byref(n_all))
    if rc:
        raise _lib.FfsError(rc, "ffs_sweep_gather", (L.ffs_comm_last_error(comm._h) or b"").decode())
    return take(allp, n_all.value), secs.value
