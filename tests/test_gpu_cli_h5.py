"""The executables on NeXus / HDF5 input and with --save-h5:
`spotfinder FILE.nxs` as the reference's tests drive it (tests/test_spotfinder.py:40-76 of the
reference) and the results_ffs.h5 table the reference's tests read back (:78-114).  Files are
produced by h5lite_tool stream2nxs (VDS master + bitshuffle-LZ4 data files) and read back with
the independent Python parser (tests/h5dissect.py)."""
import json
import re
import subprocess

import numpy as np
import pytest

import h5dissect
from ffs_b200 import synth
from oracle import oracle
from test_gpu_cli import BIN, run, write_stream

pytestmark = pytest.mark.gpu
TOOL = BIN / "h5lite_tool"
DET = json.dumps({"pixel_size_x": 0.075, "pixel_size_y": 0.075, "beam_center_x": 11.0, "beam_center_y": 7.5,
                  "distance": 150.0, "image_size_x": 300, "detector": "Eiger"})


def to_nxs(stream_dir, prefix, per_file=0):
    subprocess.run([str(TOOL), "stream2nxs", str(stream_dir), str(prefix), str(per_file)], check=True,
                   capture_output=True, timeout=120)
    return str(prefix) + ".nxs"


def expected_stills(frames, mask):
    out = []
    for f in frames:
        strong = oracle.dispersion(f, mask)
        _, spots = oracle.connected_components(strong, f, want_labels=False)
        kept, _, _ = oracle.filter_spots(spots, 3, 2.0)
        out.append((int(strong.sum()), int((spots["num_pixels"] >= 3).sum()), kept))
    return out


@pytest.mark.parametrize("exe,dtype,bits", [("spotfinder", np.uint16, 16), ("spotfinder32", np.uint32, 32)])
def test_nxs_input_same_results_as_the_stream(tmp_path, exe, dtype, bits):
    g = synth.small_geometry(300, 200)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 7, dtype, config=8, n_spots=12)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, bits)
    nxs = to_nxs(sd, tmp_path / "scan", per_file=3)  # three data files behind one VDS
    rc, lines, out, err = run(exe, [nxs, "--threads", "3", "--output-for-index"])
    assert rc == 0 and err == "", out + err      # the reference's tests require an empty stderr (test_spotfinder.py:49)
    assert "Total data size: 7x200x300" in out and "Mask has 60000 elements" in out
    assert sorted(l["file-number"] for l in lines) == list(range(7))
    want = expected_stills(frames, mask)
    for l in lines:
        strong, n_spots, kept = want[l["file-number"]]
        assert l["file"] == nxs
        assert l["num_strong_pixels"] == strong and l["n_spots_total"] == n_spots
        coms = np.stack([kept["com_x"], kept["com_y"], kept["com_z"]], axis=1).astype(np.float64).ravel()
        assert np.allclose(l["spot_centers"], coms, rtol=1e-5, atol=1e-4)
    # geometry and wavelength come from the NXmx metadata when --detector / --wavelength are absent
    assert "Got wavelength from file: 0.976" in out
    assert re.search(r"Distance:\s+150\.0 mm", out)
    # the same frames through the stream directory give the same lines
    rc2, lines2, _, _ = run(exe, [str(sd), "--threads", "3", "--output-for-index"])
    assert rc2 == 0
    strip = lambda ls: sorted((l["file-number"], l["n_spots_total"], l["num_strong_pixels"], tuple(l["spot_centers"])) for l in ls)
    assert strip(lines) == strip(lines2)


def test_nxs_wrong_bit_depth_exit_code(tmp_path):
    g = synth.small_geometry(96, 64)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 2, np.uint32, config=8, n_spots=3)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, 32)
    nxs = to_nxs(sd, tmp_path / "scan")
    rc, _, out, _ = run("spotfinder", [nxs])
    assert rc == 32 and "Data type mismatch" in out   # the service re-runs with spotfinder32 (service.py:480-492)


def test_not_an_hdf5_file_is_an_error(tmp_path):
    p = tmp_path / "junk.nxs"
    p.write_bytes(b"this is not HDF5" * 100)
    rc, _, out, _ = run("spotfinder", [str(p), "--timeout", "0.3"])
    assert rc == 1 and "Waited too long" in out


def test_uncompressed_dataset_is_read_as_pixels(tmp_path):
    """Data files written without a compression filter (`stream2nxs ... raw`): the frame loop takes
    pixels instead of chunks (Reader::chunks_are_raw_pixels / read_pixels)."""
    g = synth.small_geometry(96, 64)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 3, np.uint16, config=8, n_spots=4)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, 16)
    r = subprocess.run([str(TOOL), "stream2nxs", str(sd), str(tmp_path / "raw"), "0", "raw"], capture_output=True, timeout=120)
    assert r.returncode == 0, r.stderr.decode()
    f = h5dissect.File(tmp_path / "raw_000001.h5")
    assert f.get("/data").filters() == []
    rc, lines, out, err = run("spotfinder", [str(tmp_path / "raw.nxs"), "--threads", "2"])
    assert rc == 0, out + err
    want = expected_stills(frames, mask)
    assert sorted((l["file-number"], l["num_strong_pixels"], l["n_spots_total"]) for l in lines) == [
        (i, want[i][0], want[i][1]) for i in range(3)]


def test_save_h5_stills(tmp_path):
    g = synth.small_geometry(300, 200)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 5, np.uint16, config=8, n_spots=12)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, 16)
    outdir = tmp_path / "out"
    outdir.mkdir()
    rc, lines, out, err = run("spotfinder", [str(sd), "--threads", "2", "--save-h5", "--detector", DET], cwd=outdir)
    assert rc == 0 and err == ""
    want = expected_stills(frames, mask)
    n = sum(len(k) for _, _, k in want)
    assert f"Successfully wrote {n} 2D reflections to HDF5 file" in out and "2D spot analysis complete" in out
    f = h5dissect.File(outdir / "results_ffs.h5")
    grp = f.get("/dials/processing/group_0")
    a = grp.attrs()
    assert a["experiment_ids"].tolist() == list(range(5)) and len(set(a["identifiers"])) == 5   # one per image
    assert sorted(grp.links()) == ["id", "xyzobs.px.value"]
    xyz = f.get("/dials/processing/group_0/xyzobs.px.value").read()
    ids = f.get("/dials/processing/group_0/id").read()
    assert xyz.dtype == np.float64 and xyz.shape == (n, 3) and ids.dtype == np.int32 and ids.shape == (n,)
    exp_ids = np.concatenate([np.full(len(k), i) for i, (_, _, k) in enumerate(want)])
    assert np.array_equal(ids, exp_ids)
    exp_xyz = np.concatenate([np.stack([k["com_x"], k["com_y"], k["com_z"]], axis=1) for _, _, k in want]).astype(np.float64)
    assert np.allclose(xyz, exp_xyz, rtol=1e-5, atol=1e-4)
    # the reference's own check of this file (tests/test_spotfinder.py:98-106): x within the image
    assert xyz[:, 0].min() >= 0 and xyz[:, 0].max() <= 300 and (xyz[:, 2] == 0.5).all()  # one plane: z + 0.5


def test_save_h5_rotation_from_nxs(tmp_path):
    g = synth.small_geometry(240, 180)
    mask = synth.make_mask(g)
    sweep = synth.make_sweep(g, mask, 8, np.uint16, n_reflections=40)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, sweep, mask, 16, omega_increment=0.1)
    nxs = to_nxs(sd, tmp_path / "sweep", per_file=5)
    outdir = tmp_path / "out"
    outdir.mkdir()
    rc, lines, out, err = run("spotfinder", [nxs, "--threads", "2", "--save-h5"], cwd=outdir)
    assert rc == 0, out + err
    assert "Oscillation:" in out and "Successfully wrote 3D reflections to HDF5 file" in out
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    labels, spots = oracle.connected_components(strong, sweep)
    kept, _, _ = oracle.filter_spots(spots, 3, 2.0)
    h, w = mask.shape
    geom = dict(distance_mm=150.0, beam_center_x_px=w / 2, beam_center_y_px=h / 2, pixel_size_x_mm=0.075,
                pixel_size_y_mm=0.075, wavelength=0.976, oscillation_width_deg=0.1)
    sb, sm, depth = oracle.kabsch_variances(labels, sweep, kept, geom)
    f = h5dissect.File(outdir / "results_ffs.h5")
    grp = f.get("/dials/processing/group_0")
    assert sorted(grp.links()) == ["id", "sigma_b_variance", "sigma_m_variance", "spot_extent_z", "xyzobs.px.value"]
    assert grp.attrs()["experiment_ids"].tolist() == [0]
    col = lambda name: f.get(f"/dials/processing/group_0/{name}").read()
    xyz = col("xyzobs.px.value")
    assert xyz.shape == (len(kept), 3)
    exp = np.stack([kept["com_x"], kept["com_y"], kept["com_z"]], axis=1).astype(np.float64)
    assert np.allclose(xyz, exp, rtol=1e-5, atol=1e-4)
    assert (col("id") == 0).all() and col("id").dtype == np.int32
    assert np.array_equal(col("spot_extent_z"), depth)
    assert np.allclose(col("sigma_b_variance"), sb, rtol=1e-4, atol=1e-12)
    assert np.allclose(col("sigma_m_variance"), sm, rtol=1e-4, atol=1e-12)


def sample_frames(dtype):
    """The six frames of h5read's --sample source (h5read.c:203-276), rebuilt with numpy; frame 5's
    pcg32 stream comes from jump-ahead doubling of the LCG (state_k = A_k * s0 + C_k mod 2^64)."""
    MOD_FAST, MOD_SLOW, GAP_FAST, GAP_SLOW, SLOW, FAST = 1028, 512, 12, 38, 4362, 4148
    mask = np.ones((SLOW, FAST), np.uint8)
    for gap in range(1, 8):
        y = gap * MOD_SLOW + (gap - 1) * GAP_SLOW
        mask[y:y + GAP_SLOW] = 0
    for gap in range(1, 4):
        x = gap * MOD_FAST + (gap - 1) * GAP_FAST
        mask[:, x:x + GAP_FAST] = 0
    frames = np.zeros((6, SLOW, FAST), dtype)
    frames[1][mask != 0] = 1
    frames[2][::42, ::42] = 100
    frames[3] = np.arange(FAST, dtype=dtype)[None, :]
    frames[4] = np.arange(SLOW, dtype=dtype)[:, None]
    n = 8 * 4 * MOD_SLOW * MOD_FAST
    mult = np.uint64(6364136223846793005)
    with np.errstate(over="ignore"):
        A, C = np.array([1], np.uint64), np.array([0], np.uint64)   # state before draw k, from s0 = 0, inc | 1 = 1
        while len(A) < n:
            An, Cn = A[-1] * mult, C[-1] * mult + np.uint64(1)      # one step past the last known
            A, C = np.concatenate([A, A * An]), np.concatenate([C, A * Cn + C])
        old = C[:n]
        xs = (((old >> np.uint64(18)) ^ old) >> np.uint64(27)).astype(np.uint32)
        rot = (old >> np.uint64(59)).astype(np.uint32)
        draws = (xs >> rot) | (xs << ((np.uint32(0) - rot) & np.uint32(31)))
    vals = (draws % np.uint32(10)).astype(dtype).reshape(8, 4, MOD_SLOW, MOD_FAST)
    for my in range(8):
        for mx in range(4):
            r0, c0 = my * (MOD_SLOW + GAP_SLOW), mx * (MOD_FAST + GAP_FAST)
            frames[5][r0:r0 + MOD_SLOW, c0:c0 + MOD_FAST] = vals[my, mx]
    return frames, mask


def test_pcg32_stream_of_the_sample_generator():
    # first outputs of pcg32 with state = 0, inc = 0 (h5read.c:189-201), stepped one at a time
    state, out = 0, []
    for _ in range(5):
        old = state
        state = (old * 6364136223846793005 + 1) & (2**64 - 1)
        xs = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = old >> 59
        out.append(((xs >> rot) | (xs << ((-rot) & 31))) & 0xFFFFFFFF)
    frames, _ = sample_frames(np.uint16)
    assert frames[5][0, :5].tolist() == [v % 10 for v in out]


@pytest.mark.parametrize("exe,dtype", [("spotfinder", np.uint16), ("spotfinder32", np.uint32)])
def test_sample_source(exe, dtype, tmp_path):
    frames, mask = sample_frames(dtype)
    rc, lines, out, err = run(exe, ["--sample", "--threads", "3"], timeout=300, cwd=tmp_path)
    assert rc == 0 and err == "", out + err
    assert sorted(l["file-number"] for l in lines) == list(range(6))
    for l in lines:
        i = l["file-number"]
        strong = oracle.dispersion(frames[i], mask)
        _, spots = oracle.connected_components(strong, frames[i], want_labels=False)
        assert l["num_strong_pixels"] == int(strong.sum()), i
        assert l["n_spots_total"] == int((spots["num_pixels"] >= 3).sum()), i
    by = {l["file-number"]: l for l in lines}
    assert by[0]["num_strong_pixels"] == 0 and by[1]["num_strong_pixels"] == 0   # flat frames have no spots
    assert by[2]["num_strong_pixels"] > 0                                        # the 100-count pixels
    # the file group is required unless the environment says otherwise (spotfinder.cc:266-282)
    rc, _, out, _ = run(exe, [])
    assert rc == 1 and "One of the arguments '--sample' or 'FILE.nxs' is required" in out


def test_data_file_that_arrives_late(tmp_path):
    """The master (VDS) is there, the second data file is not: the frame loop polls
    is_image_available (spotfinder.cc:770-787) and carries on when the file lands, like a reader
    following a writer; a file that never comes ends the run with the timeout message, exit 0."""
    import threading
    import time
    g = synth.small_geometry(300, 200)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 6, np.uint16, config=8, n_spots=10)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, 16)
    nxs = to_nxs(sd, tmp_path / "scan", per_file=3)
    late = tmp_path / "scan_000002.h5"
    parked = tmp_path / "parked.h5"
    late.rename(parked)

    def deliver():
        time.sleep(1.5)
        half = parked.read_bytes()
        late.write_bytes(half[:len(half) // 3])      # first a truncated file (index not there yet) ...
        time.sleep(0.5)
        late.write_bytes(half)                        # ... then all of it
    t = threading.Thread(target=deliver)
    t.start()
    rc, lines, out, err = run("spotfinder", [nxs, "--threads", "2", "--timeout", "20"])
    t.join()
    assert rc == 0, out + err
    assert sorted(l["file-number"] for l in lines) == list(range(6))
    want = expected_stills(frames, mask)
    assert {l["file-number"]: l["num_strong_pixels"] for l in lines} == {i: want[i][0] for i in range(6)}
    waited = float(re.search(r"Total time waiting for images to appear: ([0-9.]+) (ms|s)", out).group(1))
    assert waited > 0

    late.unlink()
    rc, lines, out, err = run("spotfinder", [nxs, "--threads", "2", "--timeout", "1"])
    assert rc == 0 and "Timeout waiting for image 3" in out
    assert sorted(l["file-number"] for l in lines) == [0, 1, 2]


def test_more_images_than_the_scan_holds(tmp_path):
    g = synth.small_geometry(96, 64)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 3, np.uint16, config=8, n_spots=3)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, 16)
    nxs = to_nxs(sd, tmp_path / "scan")
    rc, lines, out, err = run("spotfinder", [nxs, "--images", "5", "--threads", "1"])
    assert rc == 1 and "Could not find data file for frame 3" in err      # h5read.c:383-386


def test_validate_script_on_a_nexus_file(tmp_path):
    """tests/validate_stream.py (the stand-in for --validate) on .nxs input: frames read through
    h5lite_tool, decoded by the oracle's bitshuffle-LZ4 port, strong pixels compared per frame."""
    import sys
    from pathlib import Path
    g = synth.small_geometry(180, 140)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 4, np.uint16, config=6, n_spots=10)
    sd = tmp_path / "stream"
    sd.mkdir()
    write_stream(sd, frames, mask, 16)
    nxs = to_nxs(sd, tmp_path / "scan", per_file=3)
    script = Path(__file__).resolve().parent / "validate_stream.py"
    for algo in ("dispersion", "dispersion_extended"):
        r = subprocess.run([sys.executable, str(script), nxs, "--algorithm", algo], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        assert r.stdout.count("identical") == 4 and "validation passed" in r.stdout
