"""Streaming API (pinned ring, async H2D, per-frame results) and 3D sweep labelling,
through the C ABI, against the CPU oracle."""
import numpy as np
import pytest

from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle
from test_gpu_cc import check_spots

pytestmark = pytest.mark.gpu


def test_streaming_matches_oracle_and_order():
    g = synth.small_geometry(600, 400)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 9, np.uint16, config=6, n_spots=15)
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=3) as sf:
        got = {}
        submitted = 0
        for i in range(len(frames)):
            if i % 2:  # both submission styles
                slot, buf = sf.acquire()
                buf[...] = frames[i]
                sf.submit(slot, 100 + i)
            else:
                sf.submit_copy(100 + i, frames[i])
            submitted += 1
            if sf.pending >= 2:
                res, spots = sf.wait()
                got[res.frame_id] = (res, spots)
        while sf.pending:
            res, spots = sf.wait()
            got[res.frame_id] = (res, spots)
        assert sorted(got) == [100 + i for i in range(len(frames))]
        for i in range(len(frames)):
            res, spots = got[100 + i]
            strong = oracle.dispersion(frames[i], mask)
            _, want = oracle.connected_components(strong, frames[i], want_labels=False)
            assert res.num_strong_pixels == int(strong.sum())
            check_spots(spots, want)
        with pytest.raises(Exception):
            sf.wait()  # nothing pending -> FFS_ERR_STATE


@pytest.mark.parametrize("row_wrap", [True, False])
def test_sweep_3d_components(row_wrap):
    g = synth.small_geometry(320, 240)
    mask = synth.make_mask(g)
    nz = 12
    sweep = synth.make_sweep(g, mask, nz, np.uint16, n_reflections=60)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    _, want = oracle.connected_components(strong, sweep, row_wrap=row_wrap, want_labels=False)
    assert (want["z_max"] > want["z_min"]).sum() > 5  # spots really span frames
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=4, cc_row_wrap=int(row_wrap)) as sf:
        # streaming path, frames submitted out of order: z = rank of the frame id
        sf.sweep_begin(nz)
        order = [3, 0, 1, 2, 7, 6, 5, 4, 8, 9, 11, 10]
        for k, i in enumerate(order):
            sf.submit_copy(1000 + 5 * i, sweep[i])
            if sf.pending >= 3:
                sf.wait()
        while sf.pending:
            sf.wait()
        got = sf.sweep_finish()
        check_spots(got, want)
        first, last = sf.sweep_export_plane(0), sf.sweep_export_plane(1)
        assert np.array_equal(np.sort(first[:, 0]), np.flatnonzero(strong[0]))
        assert np.array_equal(np.sort(last[:, 0]), np.flatnonzero(strong[-1]))
        # slot path
        sf.sweep_begin(nz)
        for i in range(nz):
            sf.upload(sweep[i], 0)
            sf.process(0, 1)
            sf.sweep_record_slot(0, i)
        got2 = sf.sweep_finish()
        check_spots(got2, want)
        kept, rs, rp = sf.filter_spots(got2, 3, 2.0)
        wk, wrs, wrp = oracle.filter_spots(want, 3, 2.0)
        assert (len(kept), rs, rp) == (len(wk), wrs, wrp)


def test_slab_sharded_sweep_merge_on_device():
    """configs[3] shape: contiguous frame slabs per rank (here: 3 slabs run one after the other
    on the one GPU), boundary planes exchanged and merged -> identical to one whole-sweep run."""
    from ffs_b200.sweep import merge_slabs, slab_bounds, slab_result_from
    g = synth.small_geometry(320, 240)
    mask = synth.make_mask(g)
    nz, world = 14, 3
    sweep = synth.make_sweep(g, mask, nz, np.uint16, n_reflections=80)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    _, want = oracle.connected_components(strong, sweep, want_labels=False)
    slabs = []
    for r in range(world):
        z0, z1 = slab_bounds(nz, world, r)
        with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=3) as sf:
            sf.sweep_begin(z1 - z0)
            for z in range(z0, z1):
                sf.submit_copy(z, sweep[z])
                if sf.pending >= 2:
                    sf.wait()
            while sf.pending:
                sf.wait()
            slabs.append(slab_result_from(sf, z0, z1 - z0))
    assert sum(len(s.spots) for s in slabs) > len(want)
    check_spots(merge_slabs(slabs, g.width, g.height), want)
