#!/usr/bin/env python
"""Executable-level throughput: the `spotfinder` binary the Zocalo service launches, fed from files,
end to end (file read -> pinned ring -> H2D of the compressed chunk -> device bitshuffle-LZ4 decode
-> threshold -> connected components -> JSON lines over the pipe), as a user of the reference
would run it.  Two sources of the same frames, both in /dev/shm:

  stream : an Eiger stream directory (start_1 / start_5 / image_NNNNNN_2, spotfinder/shmread.cc)
  nxs    : the NeXus layout H5Read opens (master with a VDS + bitshuffle-LZ4 data files),
           written by `h5lite_tool stream2nxs`

Reports the executable's own summary line (`N images in T s (G GBps) (F fps)`, spotfinder.cc:1312-1322)
per source and thread count, one JSON object on stdout.  Synthetic Eiger2 16M u16 frames; N frames
cycle over 8 distinct ones (one file per frame in the stream directory, the data files hold every frame).
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "fast-feedback-service_b200"))
sys.path.insert(0, str(ROOT / "tests"))

from ffs_b200 import bslz4_encode, synth  # noqa: E402

BIN = ROOT / "fast-feedback-service_b200" / "ffs_b200"


def run(exe, args, workdir):
    """stdout and stderr go to files: the executable prints a few lines per image, which would fill a
    pipe nobody drains while this side is still reading the JSON lines."""
    r, w = os.pipe()
    t0 = time.perf_counter()
    with open(workdir / "stdout.txt", "wb") as fo, open(workdir / "stderr.txt", "wb") as fe:
        proc = subprocess.Popen([str(BIN / exe), *args, "--pipe_fd", str(w)], pass_fds=[w], stdout=fo, stderr=fe,
                                env=dict(os.environ, FFS_DEV_KNOBS="1", FFS_CLI_TIMING="1"))
        os.close(w)
        n_lines = 0
        with os.fdopen(r) as f:
            for _ in f:
                n_lines += 1
        try:
            proc.wait(timeout=300)
        except subprocess.TimeoutExpired:
            proc.kill()
            proc.wait()
    wall = time.perf_counter() - t0
    out = (workdir / "stdout.txt").read_text(errors="replace")
    err = (workdir / "stderr.txt").read_text(errors="replace")
    wt = re.search(r"worker time.*", out)
    m = re.search(r"(\d+) images in ([0-9.]+) s \(([0-9.]+) GBps\) \(([0-9.]+) fps\)", out)
    return {"rc": proc.returncode, "json_lines": n_lines, "images": int(m.group(1)) if m else 0,
            "seconds": float(m.group(2)) if m else None, "fps": float(m.group(4)) if m else None,
            "wall_seconds_with_startup": wall, "worker_time": wt.group(0) if wt else None, "stderr": err[-300:]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=512)
    ap.add_argument("--threads", type=int, nargs="+", default=[8, 16, 32])
    ap.add_argument("--geometry", default="eiger2_16m")
    ap.add_argument("--links", action="store_true",
                    help="stream directory: hard links to 8 files instead of one file per frame (less /dev/shm, but "
                         "every reader thread then hammers the same 8 inodes)")
    ap.add_argument("--sources", nargs="+", default=["stream", "nxs"], choices=["stream", "nxs"])
    args = ap.parse_args()
    g = synth.GEOMETRIES[args.geometry]
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 8, np.uint16, config=2, n_hot=0)
    with ThreadPoolExecutor(8) as ex:
        chunks = list(ex.map(bslz4_encode.compress_chunk, frames))
    base = Path(tempfile.mkdtemp(prefix="ffs_cli_", dir="/dev/shm"))
    out = {"what": "spotfinder executable, files in /dev/shm -> JSON lines; its own fps line", "geometry": g.name,
           "frames": args.frames, "distinct_frames": 8, "mean_chunk_bytes": int(np.mean([len(c) for c in chunks])),
           "runs": []}
    try:
        sd = base / "stream"
        sd.mkdir()
        hdr = {"nimages": args.frames, "ntrigger": 1, "y_pixels_in_detector": g.height, "x_pixels_in_detector": g.width,
               "bit_depth_image": 16, "countrate_correction_count_cutoff": 60000, "wavelength": 0.976,
               "detector_distance": 150.0, "y_pixel_size": 7.5e-5, "x_pixel_size": 7.5e-5,
               "beam_center_x": g.width / 2, "beam_center_y": g.height / 2}
        (sd / "start_1").write_text(json.dumps(hdr))
        (sd / "start_4").write_bytes(b"")
        (sd / "start_5").write_bytes((mask == 0).astype(np.int32).tobytes())
        for k, c in enumerate(chunks):
            (sd / f"image_{k:06d}_2").write_bytes(c)
        for i in range(8, args.frames):
            if args.links:
                os.link(sd / f"image_{i % 8:06d}_2", sd / f"image_{i:06d}_2")
            else:
                (sd / f"image_{i:06d}_2").write_bytes(chunks[i % 8])
        if "nxs" in args.sources:
            t0 = time.perf_counter()
            subprocess.run([str(BIN / "h5lite_tool"), "stream2nxs", str(sd), str(base / "scan"), "128"], check=True,
                           capture_output=True)
            out["stream2nxs_seconds"] = time.perf_counter() - t0
        det = json.dumps({"pixel_size_x": 0.075, "pixel_size_y": 0.075, "beam_center_x": 155.0, "beam_center_y": 163.0,
                          "distance": 150.0})
        for source, path in (("stream", str(sd)), ("nxs", str(base / "scan.nxs"))):
            if source not in args.sources:
                continue
            for t in args.threads:
                r = run("spotfinder", [path, "--threads", str(t), "--detector", det], base)
                r.update(source=source, threads=t)
                print(source, t, r["fps"], r["rc"], r["worker_time"], file=sys.stderr)
                out["runs"].append(r)
    finally:
        shutil.rmtree(base, ignore_errors=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
