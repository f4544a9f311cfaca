"""BASELINE.json configs[3] against the ORACLE (not against itself):

  * a rotation sweep at Eiger2 4M size -- 64 frames, > 10^4 three-dimensional components, frames with
    more than 4096 spots (beyond the asynchronous result window of ffs_wait) -- labelled on the
    device in one go and as 3 slabs merged by the product's native merge, every integer field
    bit-exact against the oracle's whole-sweep find_3d_components restatement
    (connected_components.cc:270-470);
  * the same merge over NCCL with one rank per GPU (ffs_sweep_merge: ncclSend/ncclRecv of boundary
    planes, ncclAllGather of boundary-component records) -- needs >= 2 GPUs, skipped below that.
"""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tools"))

from ffs_b200 import synth
from ffs_b200.lib import SPOT_DTYPE
from ffs_b200.spotfinder import Spotfinder, device_count
from ffs_b200.sweep import merge_slabs_native, slab_bounds, slab_result_from
from oracle import oracle

pytestmark = pytest.mark.gpu

INT_FIELDS = ["label", "x_min", "x_max", "y_min", "y_max", "z_min", "z_max", "num_pixels", "peak_x", "peak_y",
              "peak_z", "peak_intensity", "sum_intensity"]


def check(got, want):
    assert len(got) == len(want), (len(got), len(want))
    for f in INT_FIELDS:
        assert np.array_equal(got[f], want[f]), f
    for f in ("com_x", "com_y", "com_z", "peak_centroid_distance"):
        np.testing.assert_allclose(got[f], want[f], rtol=1e-5, err_msg=f)


def stream_slab(sf, frames, z0):
    """Frames z0.. through the streaming API with sweep recording on; returns per-frame results."""
    sf.sweep_begin(len(frames))
    results = {}
    for i, f in enumerate(frames):
        sf.submit_copy(z0 + i, f)
        if sf.pending >= sf.n_slots - 1:
            res, spots = sf.wait(8192)
            results[res.frame_id] = (res, spots)
    while sf.pending:
        res, spots = sf.wait(8192)
        results[res.frame_id] = (res, spots)
    return results


def test_sweep_at_scale_single_and_three_slabs_vs_oracle():
    from bench_sweep import SweepSource
    g = synth.EIGER2_4M
    mask = synth.make_mask(g)
    nz = 64
    src = SweepSource(g, mask, nz, np.uint16, n_bg=4, n_refl=110_000)
    sweep = np.empty((nz, g.height, g.width), np.uint16)
    for z in range(nz):
        src.frame_into(z, sweep[z])
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    _, want = oracle.connected_components(strong, sweep, want_labels=False)
    assert len(want) > 10_000, len(want)

    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=4) as sf:
        # whole sweep on one GPU
        per_frame = stream_slab(sf, sweep, 0)
        whole = sf.sweep_finish()
        check(whole, want)
        # the 2D results of the same pass: more than 4096 spots in a frame (the async window)
        most = max(r.n_components for r, _ in per_frame.values())
        assert most > 4096, most
        for z in (0, nz // 2, nz - 1):
            res, spots = per_frame[z]
            _, w2 = oracle.connected_components(strong[z], sweep[z], want_labels=False)
            assert res.status == 0 and res.num_strong_pixels == int(strong[z].sum())
            assert res.n_components == len(w2)
            assert np.array_equal(spots["label"], w2["label"][:len(spots)])
            assert np.array_equal(spots["num_pixels"], w2["num_pixels"][:len(spots)])
        # three slabs, merged natively
        slabs = []
        for r in range(3):
            z0, z1 = slab_bounds(nz, 3, r)
            stream_slab(sf, sweep[z0:z1], z0)
            slabs.append(slab_result_from(sf, z0, z1 - z0))
        assert sum(len(s.spots) for s in slabs) > len(want) + 100  # the cuts really split components
        merged, sums, moments = merge_slabs_native(slabs, g.width, g.height, with_sums=True)
        check(merged, want)
        # merged exact sums / moments equal the single-GPU run's
        stream_slab(sf, sweep, 0)
        again = sf.sweep_finish()
        check(again, want)
        assert np.array_equal(sums, sf.sweep_export_sums())
        assert np.array_equal(moments, sf.sweep_export_moments())


def _nccl_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from ffs_b200.sweep import NcclComm, merge_native
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    g = synth.small_geometry(640, 480)
    mask = synth.make_mask(g)
    nz = 6 * world + 1
    sweep = synth.make_sweep(g, mask, nz, np.uint16, n_reflections=60 * world)
    z0, z1 = slab_bounds(nz, world, rank)
    comm = NcclComm.from_torch(rank)
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=3, device=rank) as sf:
        stream_slab(sf, sweep[z0:z1], z0)
        local = sf.sweep_finish()
        merged, secs = merge_native(sf, comm, z0, z1 - z0, local, gather=True)
        mine, _ = None, None
        sums = sf.sweep_export_sums()
    comm.close()
    if rank == 0:
        q.put((merged, secs, len(local), len(sums)))
    dist.barrier()
    dist.destroy_process_group()


def test_nccl_slab_merge_one_rank_per_gpu():
    world = min(device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    merged, secs, n_local, n_sums = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    g = synth.small_geometry(640, 480)
    mask = synth.make_mask(g)
    nz = 6 * world + 1
    sweep = synth.make_sweep(g, mask, nz, np.uint16, n_reflections=60 * world)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    _, want = oracle.connected_components(strong, sweep, want_labels=False)
    got = np.zeros(len(merged), SPOT_DTYPE)
    got[...] = merged
    check(got, want)
    assert 0 < secs < 5.0
