"""CPU-only: slab-boundary merging of 3D components (ffs_b200/sweep.py), single-process and
over torch.distributed with the gloo backend (world_size 2), against the oracle's
whole-sweep labelling.  Slab inputs are built from the oracle, so no GPU is involved."""
import os
import socket

import numpy as np
import pytest

from ffs_b200 import synth
from ffs_b200.lib import SPOT_DTYPE
from ffs_b200.sweep import SlabResult, merge_slabs, slab_bounds
from oracle import oracle

INT_FIELDS = ["label", "x_min", "x_max", "y_min", "y_max", "z_min", "z_max", "num_pixels", "peak_x", "peak_y",
              "peak_z", "peak_intensity", "sum_intensity"]


def make_sweep(nz=10, w=160, h=120, seed_refl=50):
    g = synth.small_geometry(w, h)
    mask = synth.make_mask(g)
    sweep = synth.make_sweep(g, mask, nz, np.uint16, n_reflections=seed_refl)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    return g, sweep, strong


def oracle_slab(strong, sweep, z0, z1) -> SlabResult:
    s, im = strong[z0:z1], sweep[z0:z1]
    labels, spots = oracle.connected_components(s, im, row_wrap=True)
    out = np.zeros(len(spots), SPOT_DTYPE)
    for f in SPOT_DTYPE.names:
        out[f] = spots[f]
    comp_of_label = {int(l): i for i, l in enumerate(spots["label"])}
    zz, yy, xx = np.nonzero(s)
    comp = np.array([comp_of_label[int(l)] for l in labels[zz, yy, xx]], dtype=np.int64)
    inten = im[zz, yy, xx].astype(object)
    sums = np.zeros((len(spots), 4), dtype=object)
    for c, i, x, y, z in zip(comp, inten, xx, yy, zz):
        sums[c] += np.array([i, (2 * int(x) + 1) * i, (2 * int(y) + 1) * i, (2 * int(z) + 1) * i], dtype=object)
    w = s.shape[2]

    def plane(zl):
        sel = zz == zl
        return np.stack([(yy[sel] * w + xx[sel]).astype(np.uint32), comp[sel].astype(np.uint32)], axis=1)

    return SlabResult(z0, z1 - z0, out, sums.astype(np.uint64), plane(0), plane(z1 - z0 - 1))


def check(got, want):
    assert len(got) == len(want)
    for f in INT_FIELDS:
        assert np.array_equal(got[f], want[f]), f
    for f in ("com_x", "com_y", "com_z", "peak_centroid_distance"):
        np.testing.assert_allclose(got[f], want[f], rtol=1e-5, err_msg=f)


@pytest.mark.parametrize("world", [1, 2, 3, 5])
def test_merge_matches_whole_sweep(world):
    g, sweep, strong = make_sweep()
    _, want = oracle.connected_components(strong, sweep, want_labels=False)
    slabs = [oracle_slab(strong, sweep, *slab_bounds(len(sweep), world, r)) for r in range(world)]
    if world > 1:  # slab cuts really split components
        assert sum(len(s.spots) for s in slabs) > len(want)
    check(merge_slabs(slabs, g.width, g.height), want)


def test_component_spanning_three_slabs_and_empty_boundary():
    g, sweep, strong = make_sweep(nz=6, seed_refl=10)
    strong[:, 40:43, 50:53] = 1          # a column of signal through every frame
    sweep[:, 40:43, 50:53] = 500
    strong[3] = 0                         # ... except an empty frame: breaks every chain there
    _, want = oracle.connected_components(strong, sweep, want_labels=False)
    slabs = [oracle_slab(strong, sweep, z, z + 1) for z in range(6)]  # one frame per "rank"
    check(merge_slabs(slabs, g.width, g.height), want)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from ffs_b200.sweep import gather_and_merge
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g, sweep, strong = make_sweep()
    local = oracle_slab(strong, sweep, *slab_bounds(len(sweep), world, rank))
    merged = gather_and_merge(local, g.width, g.height)
    if rank == 0:
        q.put(merged)
    dist.barrier()
    dist.destroy_process_group()


def test_gather_and_merge_gloo_world2():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    g, sweep, strong = make_sweep()
    _, want = oracle.connected_components(strong, sweep, want_labels=False)
    check(merged, want)
