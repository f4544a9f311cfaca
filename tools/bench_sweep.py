"""BASELINE.json configs[3]: 3D connected components over a synthetic rotation sweep, contiguous
frame slabs per GPU, boundary-plane label merge over torch.distributed (NCCL on GPUs).

    python tools/bench_sweep.py --frames 360                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29533 tools/bench_sweep.py --frames 3600      # 8 slabs

Every rank synthesises ONLY its slab (8 Poisson background frames cycled, the sweep's
reflections stamped on top with a Gaussian rocking curve), streams it through the pinned ring
with sweep recording on, labels the slab on its GPU (ffs_sweep_finish), and the slabs are merged
natively over NCCL (ffs_sweep_merge: boundary planes by ncclSend/ncclRecv, boundary-component
records by ncclAllGather, csrc/api_multi.inl).  Prints one
JSON line: frames, spots, and the time of each phase (max over ranks).  Host-side frame
synthesis is inside `stream_s` (it is the slow part of that phase and is reported separately as
`synth_s`); `finish_s` + `merge_s` are the 3D labelling itself.  --check re-runs the whole sweep
on rank 0 alone and compares (use a size that fits one GPU's patience).
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "fast-feedback-service_b200"))

import numpy as np
import torch
import torch.distributed as dist

from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from ffs_b200.sweep import NcclComm, merge_native, merge_slabs_native, slab_bounds, slab_result_from


class SweepSource:
    """Deterministic frames of a rotation sweep, any frame on demand."""

    def __init__(self, geom, mask, n_frames, dtype, n_bg=8, n_refl=None):
        self.g, self.mask, self.n, self.dtype = geom, mask, n_frames, dtype
        self.bg = [synth.make_frame(geom, mask, dtype, 4, k, background=2.0, n_hot=0, n_spots=0) for k in range(n_bg)]
        rng = np.random.Generator(np.random.PCG64(424242))
        if n_refl is None:
            n_refl = max(8, int(round(40000 * (n_frames / 3600) * geom.pixels / synth.EIGER2_16M.pixels)))
        self.rx = rng.uniform(8, geom.width - 8, n_refl)
        self.ry = rng.uniform(8, geom.height - 8, n_refl)
        self.rz = rng.uniform(0, n_frames, n_refl)
        self.amp = np.exp(rng.uniform(np.log(50.0), np.log(3000.0), n_refl))
        self.sig = rng.uniform(0.8, 1.8, n_refl)
        self.sz = rng.uniform(0.3, 1.5, n_refl)
        order = np.argsort(self.rz)
        for name in ("rx", "ry", "rz", "amp", "sig", "sz"):
            setattr(self, name, getattr(self, name)[order])

    def frame_into(self, z, out):
        np.copyto(out, self.bg[z % len(self.bg)])
        lo, hi = np.searchsorted(self.rz, [z + 0.5 - 6.0, z + 0.5 + 6.0])
        for k in range(lo, hi):
            wgt = np.exp(-0.5 * ((z + 0.5 - self.rz[k]) / self.sz[k]) ** 2)
            if wgt < 1e-3:
                continue
            x0, y0, s = self.rx[k], self.ry[k], self.sig[k]
            xa, ya = int(x0) - 5, int(y0) - 5
            xs = np.arange(xa, xa + 11) + 0.5
            ys = np.arange(ya, ya + 11) + 0.5
            patch = self.amp[k] * wgt * np.exp(-0.5 * (((xs[None, :] - x0) / s) ** 2 + ((ys[:, None] - y0) / s) ** 2))
            add = np.floor(patch).astype(np.int64)
            sub = out[ya:ya + 11, xa:xa + 11]
            valid = self.mask[ya:ya + 11, xa:xa + 11] != 0
            sub[...] = np.where(valid, np.minimum(sub.astype(np.int64) + add, 60000), sub).astype(self.dtype)


def run_slab(sf, src, z0, z1, times):
    sf.sweep_begin(z1 - z0)
    pending = 0
    t_syn = 0.0
    t0 = time.perf_counter()
    for z in range(z0, z1):
        slot, buf = sf.acquire()
        ts = time.perf_counter()
        src.frame_into(z, buf)
        t_syn += time.perf_counter() - ts
        sf.submit(slot, z)
        pending += 1
        if pending >= sf.n_slots - 1:
            sf.wait()
            pending -= 1
    while pending:
        sf.wait()
        pending -= 1
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    slab = slab_result_from(sf, z0, z1 - z0)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    times.update(stream_s=t1 - t0, synth_s=t_syn, finish_s=t2 - t1)
    return slab


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=360)
    ap.add_argument("--geom", default="eiger2_16m")
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    geom = synth.GEOMETRIES[a.geom]
    mask = synth.make_mask(geom)
    src = SweepSource(geom, mask, a.frames, np.uint16)
    z0, z1 = slab_bounds(a.frames, world, rank)
    times = {}
    comm = NcclComm.from_torch(local) if world > 1 else None
    with Spotfinder(geom.width, geom.height, np.uint16, mask, n_slots=6, device=local) as sf:
        slab = run_slab(sf, src, z0, z1, times)
        if world > 1:
            dist.barrier()
        tm = time.perf_counter()
        if comm is not None:
            merged, times["exchange_s"] = merge_native(sf, comm, z0, z1 - z0, slab.spots, gather=True)
        else:
            merged, times["exchange_s"] = merge_slabs_native([slab], geom.width, geom.height), 0.0
        times["merge_s"] = time.perf_counter() - tm
        kept, rs, rp = sf.filter_spots(merged, 3, 2.0)
        ok = None
        if a.check and rank == 0:
            t = {}
            whole = run_slab(sf, src, 0, a.frames, t)
            ref = merge_slabs_native([whole], geom.width, geom.height)
            ok = bool(len(ref) == len(merged) and np.array_equal(ref["label"], merged["label"])
                      and np.array_equal(ref["num_pixels"], merged["num_pixels"])
                      and np.array_equal(ref["sum_intensity"], merged["sum_intensity"])
                      and np.allclose(ref["com_z"], merged["com_z"], rtol=1e-5))
    vals = torch.tensor([times["stream_s"], times["synth_s"], times["finish_s"], times["merge_s"],
                         float(len(slab.spots))], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = vals.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = vals.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    else:
        mx, sm = vals, vals
    if rank == 0:
        print(json.dumps({
            "workload": f"{a.geom} u16 rotation sweep, dispersion + 3D connected components, slab per GPU",
            "frames": a.frames, "n_gpus": world, "slab_components_total": int(sm[4].item()),
            "components_after_merge": int(len(merged)), "spots_after_filters": int(len(kept)),
            "stream_s": round(mx[0].item(), 3), "synth_s_inside_stream": round(mx[1].item(), 3),
            "finish_3d_s": round(mx[2].item(), 3), "merge_s": round(mx[3].item(), 4),
            "merge": "native: ncclSend/ncclRecv of boundary planes + ncclAllGather of boundary-component records",
            "labelling_frames_per_s": round(a.frames / (mx[2].item() + mx[3].item()), 1),
            "identical_to_single_gpu": ok}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
