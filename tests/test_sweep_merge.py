"""CPU-only: slab-boundary merging of 3D components (ffs_b200/sweep.py), single-process and
over torch.distributed with the gloo backend (world_size 2), against the oracle's
whole-sweep labelling.  Slab inputs are built from the oracle, so no GPU is involved."""
import os
import socket

import numpy as np
import pytest

from ffs_b200 import synth
from ffs_b200.lib import SPOT_DTYPE
from ffs_b200.sweep import SlabResult, merge_slabs, merge_slabs_native, slab_bounds
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
    moments = np.zeros((len(spots), 4), dtype=object)
    for c, i, x, y, z in zip(comp, inten, xx, yy, zz):
        u, v, q = int(x) - int(out["x_min"][c]), int(y) - int(out["y_min"][c]), int(z) - int(out["z_min"][c])
        moments[c] += np.array([i * u * u, i * v * v, i * u * v, i * q * q], dtype=object)
    w = s.shape[2]

    def plane(zl):
        sel = zz == zl
        return np.stack([(yy[sel] * w + xx[sel]).astype(np.uint32), comp[sel].astype(np.uint32)], axis=1)

    return SlabResult(z0, z1 - z0, out, sums.astype(np.uint64), plane(0), plane(z1 - z0 - 1),
                      moments.astype(np.uint64))


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
    # the product's native merge (csrc/api_multi.inl, ffs_sweep_merge_host: plain host code, no GPU)
    native = merge_slabs_native(slabs, g.width, g.height)
    check(native, want)
    for f in ("com_x", "com_y", "com_z", "peak_centroid_distance"):  # same float operations as the device
        assert np.array_equal(native[f], merge_slabs(slabs, g.width, g.height)[f], equal_nan=True), f


@pytest.mark.parametrize("world", [2, 4])
def test_merged_sums_and_moments_are_exact(world):
    """Second moments (the inputs of the Kabsch-space variances) survive slab merging exactly."""
    g, sweep, strong = make_sweep(nz=9)
    whole = oracle_slab(strong, sweep, 0, len(sweep))
    slabs = [oracle_slab(strong, sweep, *slab_bounds(len(sweep), world, r)) for r in range(world)]
    spots, sums, moments = merge_slabs(slabs, g.width, g.height, with_sums=True)
    check(spots, whole.spots)
    assert np.array_equal(sums, whole.sums)
    assert np.array_equal(moments, whole.moments)
    nspots, nsums, nmom = merge_slabs_native(slabs, g.width, g.height, with_sums=True)
    check(nspots, whole.spots)
    assert np.array_equal(nsums, whole.sums)
    assert np.array_equal(nmom, whole.moments)


def test_oracle_kabsch_variances_known_answers():
    """Hand-checkable cases of the per-pixel restatement: a one-pixel spot has zero variance;
    two equal pixels one frame apart have var_e3 = (half a frame of rotation * zeta)^2."""
    strong = np.zeros((3, 40, 50), np.uint8)
    image = np.zeros((3, 40, 50), np.uint16)
    strong[0, 5, 7] = 1
    image[0, 5, 7] = 9
    strong[1:3, 30, 20] = 1
    image[1:3, 30, 20] = 100
    labels, spots = oracle.connected_components(strong, image)
    geom = dict(distance_mm=150.0, beam_center_x_px=25.0, beam_center_y_px=20.0, pixel_size_x_mm=0.075,
                pixel_size_y_mm=0.075, wavelength=1.0, oscillation_width_deg=0.1)
    sb, sm, depth = oracle.kabsch_variances(labels, image, spots, geom)
    assert list(depth) == [1, 2]
    assert sb[0] == 0.0 and sm[0] == 0.0
    assert abs(sb[1]) < 1e-30
    # zeta = e1.x with e1 = normalise(s1 x s0), s1 = lab coordinate of the centre of mass
    s1 = np.array([(20.5 - 25.0) * 0.075, -(30.5 - 20.0) * 0.075, -150.0])
    e1 = np.cross(s1, np.array([0, 0, -1.0]))
    zeta = (e1 / np.linalg.norm(e1))[0]
    np.testing.assert_allclose(sm[1], (0.5 * 0.1 * np.pi / 180.0 * zeta) ** 2, rtol=1e-12)


def test_component_spanning_three_slabs_and_empty_boundary():
    g, sweep, strong = make_sweep(nz=6, seed_refl=10)
    strong[:, 40:43, 50:53] = 1          # a column of signal through every frame
    sweep[:, 40:43, 50:53] = 500
    strong[3] = 0                         # ... except an empty frame: breaks every chain there
    _, want = oracle.connected_components(strong, sweep, want_labels=False)
    slabs = [oracle_slab(strong, sweep, z, z + 1) for z in range(6)]  # one frame per "rank"
    check(merge_slabs(slabs, g.width, g.height), want)
    check(merge_slabs_native(slabs, g.width, g.height), want)
    check(merge_slabs_native(slabs[::-1], g.width, g.height), want)  # slab order is by z0, not by position
    # non-adjacent slabs (a missing frame between them) must not be joined
    gap = [oracle_slab(strong, sweep, 0, 2), oracle_slab(strong, sweep, 4, 6)]
    assert len(merge_slabs_native(gap, g.width, g.height)) == len(gap[0].spots) + len(gap[1].spots)


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
