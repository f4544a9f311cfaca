"""The spotfinder / spotfinder32 executables as the Zocalo service drives them
(src/ffs/service.py:419-507): stream directory in, JSON lines over an inherited pipe fd out,
stdout line formats (tests/test_spotfinder.py:26-29 of the reference), exit codes."""
import json
import os
import re
import struct
import subprocess
from pathlib import Path

import numpy as np
import pytest

from bslz4_ref import compress_chunk
from ffs_b200 import synth
from oracle import oracle

pytestmark = pytest.mark.gpu
BIN = Path(__file__).resolve().parents[1] / "fast-feedback-service_b200" / "ffs_b200"


def write_stream(d: Path, frames, mask, bits, omega_increment=None, start=0):
    h, w = mask.shape
    hdr = {"nimages": len(frames), "ntrigger": 1, "y_pixels_in_detector": h, "x_pixels_in_detector": w,
           "bit_depth_image": bits, "countrate_correction_count_cutoff": 60000, "wavelength": 0.976,
           "detector_distance": 150.0, "y_pixel_size": 7.5e-5, "x_pixel_size": 7.5e-5,
           "beam_center_x": w / 2, "beam_center_y": h / 2}
    if omega_increment is not None:
        hdr.update(omega_start=0.0, omega_increment=omega_increment)
    (d / "start_1").write_text(json.dumps(hdr))
    (d / "start_4").write_bytes(b"")
    (d / "start_5").write_bytes((mask == 0).astype(np.int32).tobytes())
    for i, f in enumerate(frames):
        (d / f"image_{start + i:06d}_2").write_bytes(compress_chunk(f))


def run(exe, args, timeout=120, host_decode=False, cwd=None):
    r, w = os.pipe()
    env = dict(os.environ, FFS_DEV_KNOBS="1", FFS_HOST_DECODE="1" if host_decode else "0")
    proc = subprocess.Popen([str(BIN / exe), *args, "--pipe_fd", str(w)], pass_fds=[w], stdout=subprocess.PIPE,
                            stderr=subprocess.PIPE, env=env, cwd=cwd)
    os.close(w)
    lines = []
    with os.fdopen(r) as f:
        for line in f:
            lines.append(json.loads(line))
    out, err = proc.communicate(timeout=timeout)
    return proc.returncode, lines, out.decode(), err.decode()


@pytest.mark.parametrize("host_decode", [False, True], ids=["device-bslz4", "host-bslz4"])
def test_stills_json_lines_and_stdout(tmp_path, host_decode):
    g = synth.small_geometry(300, 200)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 5, np.uint16, config=8, n_spots=12)
    write_stream(tmp_path, frames, mask, 16)
    rc, lines, out, err = run("spotfinder", [str(tmp_path), "--threads", "3", "--output-for-index",
                                              "--detector", json.dumps({"pixel_size_x": 0.075, "pixel_size_y": 0.075,
                                                                        "beam_center_x": 11.0, "beam_center_y": 7.5,
                                                                        "distance": 150.0, "image_size_x": 300,
                                                                        "detector": "Eiger"})],
                               host_decode=host_decode)
    assert rc == 0 and err == ""
    assert sorted(l["file-number"] for l in lines) == list(range(5))
    found = {int(m.group(1)): int(m.group(2)) for m in re.finditer(r"image\s+(\d+).*?(\d+)\s+strong pixels", out)}
    for l in lines:
        i = l["file-number"]
        assert list(l) == ["file", "file-number", "n_spots_total", "num_strong_pixels", "spot_centers"]
        assert l["file"] == str(tmp_path)
        strong = oracle.dispersion(frames[i], mask)
        _, spots = oracle.connected_components(strong, frames[i], want_labels=False)
        kept, _, _ = oracle.filter_spots(spots, 3, 2.0)
        assert l["num_strong_pixels"] == int(strong.sum()) == found[i]
        assert l["n_spots_total"] == int((spots["num_pixels"] >= 3).sum())
        want = np.stack([kept["com_x"], kept["com_y"], kept["com_z"]], axis=1).astype(np.float64).ravel()
        np.testing.assert_allclose(np.array(l["spot_centers"]), want, rtol=1e-5)
    assert re.search(r"\n5 images in [\d.]+ s \([\d.]+ GBps\) \([\d.]+ fps\)", out)


def test_trusted_range_from_the_stream_header(tmp_path):
    """countrate_correction_count_cutoff -> get_trusted_range()[1] -> the kernels
    (spotfinder.cc:482,868,879): unmasked pixels above it are invalid."""
    g = synth.small_geometry(300, 200)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 3, np.uint16, config=8, n_spots=12)
    rng = np.random.default_rng(4)
    for f in frames:
        f[rng.integers(0, 200, 12), rng.integers(0, 300, 12)] = 60001 + rng.integers(0, 5000, 12)
    write_stream(tmp_path, frames, mask, 16)  # cutoff 60000
    rc, lines, out, err = run("spotfinder", [str(tmp_path), "--threads", "2"])
    assert rc == 0 and err == ""
    changed = 0
    for l in lines:
        f = frames[l["file-number"]]
        folded = (mask.astype(bool) & (f <= 60000)).astype(np.uint8)
        want = oracle.dispersion(f, folded)
        changed += int(want.sum()) != int(oracle.dispersion(f, mask).sum())
        _, spots = oracle.connected_components(want, f, want_labels=False)
        assert l["num_strong_pixels"] == int(want.sum())
        assert l["n_spots_total"] == int((spots["num_pixels"] >= 3).sum())
    assert len(lines) == 3 and changed > 0


def test_bit_depth_exit_code_and_spotfinder32(tmp_path):
    g = synth.small_geometry(200, 150)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 2, np.uint32, config=5, n_spots=8)
    write_stream(tmp_path, frames, mask, 32)
    rc, lines, out, _ = run("spotfinder", [str(tmp_path)])
    assert rc == 32 and not lines and "Data type mismatch" in out  # service relaunches spotfinder32 on 32
    rc, lines, out, err = run("spotfinder32", [str(tmp_path), "-a", "dispersion_extended"])
    assert rc == 0 and err == ""
    for l in lines:
        want = oracle.dispersion_extended(frames[l["file-number"]], mask)
        assert l["num_strong_pixels"] == int(want.sum())


def test_rotation_sweep_3d_and_start_index(tmp_path):
    g = synth.small_geometry(240, 180)
    mask = synth.make_mask(g)
    sweep = synth.make_sweep(g, mask, 8, np.uint16, n_reflections=40)
    write_stream(tmp_path, sweep, mask, 16, omega_increment=0.1)
    outdir = tmp_path / "out"
    outdir.mkdir()
    rc, lines, out, err = run("spotfinder", [str(tmp_path), "--threads", "2", "--writeout", "--start-index", "2",
                                              "--images", "6"], cwd=outdir)
    assert rc == 0 and err == ""
    assert sorted(l["file-number"] for l in lines) == list(range(6))  # 0-based loop counter (N16)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep[2:]])
    labels, spots = oracle.connected_components(strong, sweep[2:])
    kept, rs, rp = oracle.filter_spots(spots, 3, 2.0)
    assert int(re.search(r"Calculated\s+(\d+)\s+spots", out).group(1)) == len(spots)
    # estimated sigma_b / sigma_m over the kept reflections (spotfinder.cc:1180-1212)
    h, w = mask.shape
    geom = dict(distance_mm=150.0, beam_center_x_px=w / 2, beam_center_y_px=h / 2, pixel_size_x_mm=0.075,
                pixel_size_y_mm=0.075, wavelength=0.976, oscillation_width_deg=0.1)
    sb, sm, depth = oracle.kabsch_variances(labels, sweep[2:], kept, geom)
    got_b = float(re.search(r"Estimated sigma_b \(degrees\): ([0-9.]+)", out).group(1))
    assert got_b == pytest.approx(np.sqrt(sb.mean()) * 180 / np.pi, rel=1e-4, abs=2e-6)
    deep = depth >= 5
    m = re.search(r"Estimated sigma_m \(degrees\): ([0-9.]+), calculated on (\d+) spots", out)
    if deep.any():
        assert int(m.group(2)) == int(deep.sum())
        assert float(m.group(1)) == pytest.approx(np.sqrt(sm[deep].mean()) * 180 / np.pi, rel=1e-4, abs=2e-6)
    else:
        assert m is None
    if rs:
        assert int(re.search(r"Filtered\s+(\d+)\s+spots with size < 3 pixels", out).group(1)) == rs
    if rp:
        assert int(re.search(r"Filtered\s+(\d+)\s+spots with peak-centroid distance > 2", out).group(1)) == rp
    assert int(re.search(r"Found (\d+) spots", out).group(1)) == len(kept)


def read_png_rgb(path):
    """Decode the stored-deflate RGB8 PNGs the executable writes (zlib inflates them like any other)."""
    import zlib
    b = path.read_bytes()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(typ + data)
        if typ == b"IHDR":
            w, h, depth, colour = struct.unpack(">IIBB", data[:10])
            assert (depth, colour) == (8, 2)
        elif typ == b"IDAT":
            idat += data
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)
    assert not raw[:, 0].any()  # filter type 0 on every row
    return raw[:, 1:].reshape(h, w, 3)


def test_writeout_files(tmp_path):
    """--writeout: mask_source.png, mask_calculated.png, image_NNNNN.png, pixels_NNNNN.txt
    (spotfinder.cc:626-679,939-994)."""
    g = synth.small_geometry(200, 160)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 2, np.uint16, config=6, n_spots=12)
    write_stream(tmp_path, frames, mask, 16)
    outdir = tmp_path / "out"
    outdir.mkdir()
    rc, lines, out, err = run("spotfinder", [str(tmp_path), "--threads", "1", "--writeout", "--dmin", "30"], cwd=outdir)
    assert rc == 0 and err == ""
    src = read_png_rgb(outdir / "mask_source.png")
    assert np.array_equal(src[..., 1] == 255, mask != 0) and (src[..., 0] == 255).all()
    calc = read_png_rgb(outdir / "mask_calculated.png")
    valid = calc[..., 1] == 255
    assert valid.sum() < (mask != 0).sum() and not (valid & (mask == 0)).any()  # resolution filter only removes
    for i, f in enumerate(frames):
        strong = oracle.dispersion(f, valid.astype(np.uint8))
        yy, xx = np.nonzero(strong)
        want_txt = "".join(f"{x:4d}, {y:4d}\n" for x, y in zip(xx, yy))
        assert (outdir / f"pixels_{i:05d}.txt").read_text() == want_txt
        img = read_png_rgb(outdir / f"image_{i:05d}.png")
        red = (img[..., 0] == 255) & (img[..., 1] == 0) & (img[..., 2] == 0)
        assert np.array_equal(red, strong != 0)
        # a pixel away from spots and boxes keeps the reference's grey ramp: max(0, 255.99 - 10 v)
        plain = ~red & ~((img[..., 2] == 255) & (img[..., 0] == 0))
        want_grey = np.maximum(0.0, np.float32(255.99) - f.astype(np.float32) * 10).astype(np.uint8)
        assert np.array_equal(img[..., 0][plain], want_grey[plain])


def test_validate_stream_script(tmp_path):
    """tests/validate_stream.py: the field parity check that stands in for --validate."""
    import sys
    g = synth.small_geometry(180, 140)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 3, np.uint16, config=6, n_spots=10)
    write_stream(tmp_path, frames, mask, 16)
    script = Path(__file__).resolve().parent / "validate_stream.py"
    for algo in ("dispersion", "dispersion_extended"):
        r = subprocess.run([sys.executable, str(script), str(tmp_path), "--algorithm", algo], capture_output=True,
                           text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        assert r.stdout.count("identical") == 3 and "validation passed" in r.stdout


def test_list_devices_and_bad_args():
    out = subprocess.run([str(BIN / "spotfinder"), "--list-devices"], capture_output=True, text=True)
    assert out.returncode == 0 and re.search(r"0: .+ \(CUDA \d+\.\d+\)", out.stdout)  # service.py:220-238 probes this
    assert subprocess.run([str(BIN / "spotfinder")], capture_output=True).returncode == 1
    assert subprocess.run([str(BIN / "spotfinder"), "/tmp", "-a", "sobel"], capture_output=True).returncode == 1


def test_single_thread_timing_block(tmp_path):
    """--threads 1 (the default) prints the reference's per-image Copy / Kernel / Post Copy block
    (spotfinder.cc:1055-1076) instead of the one-line multi-thread report."""
    g = synth.small_geometry(300, 200)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 3, np.uint16, config=8, n_spots=12)
    write_stream(tmp_path, frames, mask, 16)
    rc, lines, out, err = run("spotfinder", [str(tmp_path)])
    assert rc == 0 and err == "" and len(lines) == 3
    plain = re.sub(r"\x1b\[[0-9;]*m", "", out)
    blocks = re.findall(r"Thread\s+0 finished image\s+(\d+)\n\s+Copy:\s+([\d.]+) ms\n\s+Kernel:\s+([\d.]+) ms\n"
                        r"\s+Post Copy:\s+([\d.]+) ms\n\s+Post:\s+([\d.]+) ms\n\s+\S+\n\s+Total:\s+([\d.]+) ms "
                        r"\(([\d.]+) GBps\)\n\s+(\d+) strong pixels\n\s+(\d+) filtered reflections \((\d+) pixels\)",
                        plain)
    assert sorted(int(b[0]) for b in blocks) == [0, 1, 2]
    for b in blocks:
        strong = oracle.dispersion(frames[int(b[0])], mask)
        assert int(b[7]) == int(strong.sum())
        assert float(b[5]) == pytest.approx(float(b[1]) + float(b[2]) + float(b[3]), abs=0.2)


def test_devices_option_shards_stills_and_rotation(tmp_path):
    """--devices: several contexts inside one process, frames sharded frame by frame (stills) or in
    contiguous slabs merged through their boundary planes (rotation).  Listing device 0 three times
    exercises the whole multi-context path on a one-GPU box; the results must not change."""
    g = synth.small_geometry(240, 180)
    mask = synth.make_mask(g)
    stills = tmp_path / "stills"
    stills.mkdir()
    frames = synth.make_frames(g, mask, 7, np.uint16, config=8, n_spots=10)
    write_stream(stills, frames, mask, 16)
    rc, lines, out, err = run("spotfinder", [str(stills), "--threads", "4", "--devices", "0,0,0"])
    assert rc == 0 and err == "" and sorted(l["file-number"] for l in lines) == list(range(7))
    for l in lines:
        strong = oracle.dispersion(frames[l["file-number"]], mask)
        assert l["num_strong_pixels"] == int(strong.sum())
    rot = tmp_path / "rot"
    rot.mkdir()
    sweep = synth.make_sweep(g, mask, 10, np.uint16, n_reflections=50)
    write_stream(rot, sweep, mask, 16, omega_increment=0.1)
    strong = np.stack([oracle.dispersion(f, mask) for f in sweep])
    _, spots = oracle.connected_components(strong, sweep, want_labels=False)
    kept, _, _ = oracle.filter_spots(spots, 3, 2.0)
    texts = {}
    for name, extra in (("one", []), ("three", ["--devices", "0-0,0,0"])):
        outdir = tmp_path / name
        outdir.mkdir()
        rc, lines, out, err = run("spotfinder", [str(rot), "--threads", "3", "--writeout", *extra], cwd=outdir)
        assert rc == 0 and err == "" and len(lines) == 10
        assert int(re.search(r"Calculated\s+(\d+)\s+spots", out).group(1)) == len(spots)
        assert int(re.search(r"Found\s+(\d+)\s+spots", out).group(1)) == len(kept)
        texts[name] = (outdir / "3d_reflections.txt").read_text()
        sig = re.search(r"Estimated sigma_b \(degrees\): ([0-9.]+)", out).group(1)
        texts[name + "_sigma"] = sig
    # the reference's text format and order (spotfinder.cc:1119-1148; tests/3d_connected_components.sh:55-64)
    assert texts["one"].splitlines() == oracle.format_reflections_3d(kept)
    assert texts["three"] == texts["one"]
    assert texts["three_sigma"] == texts["one_sigma"]
