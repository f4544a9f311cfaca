"""Field parity check for a stream directory -- the role of the reference's `--validate`
(spotfinder.cc:1012-1052: re-run the CPU algorithm on each image and compare the strong pixels).

The GPU executable contains no CPU implementation (the product path must fail rather than fall
back), so the comparison lives here, in the test tree: the unmodified `spotfinder` executable
is run with --writeout on the directory, and every pixels_NNNNN.txt it writes is compared with
the CPU oracle's DIALS dispersion of the same decoded frame.

    python tests/validate_stream.py <stream dir> [--images N] [--algorithm dispersion|dispersion_extended]

Exit code 0 = every frame's strong-pixel set is identical.  Needs a GPU (runs the executable).
"""
import argparse
import json
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "fast-feedback-service_b200"))
sys.path.insert(0, str(ROOT / "tests"))

from ffs_b200 import lib  # noqa: E402
from oracle import oracle  # noqa: E402


def decompress_chunk(chunk: bytes, dtype, n: int) -> np.ndarray:
    """Host bitshuffle-LZ4 decoder of the executable (csrc/cli/bslz4.hpp), exported for the tests."""
    out = np.zeros(n, dtype)
    rc = lib.load().ffs_bslz4_decode_chunk(chunk, len(chunk), out.ctypes.data, n, out.dtype.itemsize)
    if rc:
        raise ValueError(f"bitshuffle-LZ4 decode failed: {rc}")
    return out


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("stream")
    ap.add_argument("--images", type=int, default=0)
    ap.add_argument("--algorithm", default="dispersion")
    a = ap.parse_args()
    d = Path(a.stream)
    hdr = json.loads((d / "start_1").read_text())
    h, w = hdr["y_pixels_in_detector"], hdr["x_pixels_in_detector"]
    bits = hdr["bit_depth_image"]
    dtype = np.uint16 if bits == 16 else np.uint32
    mask = (np.frombuffer((d / "start_5").read_bytes(), np.int32).reshape(h, w) == 0).astype(np.uint8)
    n = a.images or hdr["nimages"]
    exe = ROOT / "fast-feedback-service_b200" / "ffs_b200" / ("spotfinder" if bits == 16 else "spotfinder32")
    with tempfile.TemporaryDirectory() as tmp:
        cmd = [str(exe), str(d), "--images", str(n), "--threads", "2", "--writeout", "--algorithm", a.algorithm]
        r = subprocess.run(cmd, cwd=tmp, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stdout[-2000:], r.stderr[-2000:])
            return r.returncode or 1
        bad = 0
        for i in range(n):
            frame = decompress_chunk((d / f"image_{i:06d}_2").read_bytes(), dtype, h * w).reshape(h, w)
            fn = oracle.dispersion if a.algorithm == "dispersion" else oracle.dispersion_extended
            # trusted range: pixels above the (pixel_t-narrowed) count cutoff are invalid
            cutoff = int(np.array(int(hdr.get("countrate_correction_count_cutoff", 0))).astype(dtype))
            m = mask & (frame <= cutoff) if cutoff > 0 else mask
            yy, xx = np.nonzero(fn(frame, m.astype(np.uint8)))
            want = "".join(f"{x:4d}, {y:4d}\n" for x, y in zip(xx, yy))
            got = (Path(tmp) / f"pixels_{i:05d}.txt").read_text()
            ok = got == want
            bad += not ok
            print(f"image {i:5d}: {len(xx):7d} strong pixels  {'identical' if ok else 'MISMATCH'}")
    print("validation", "passed" if bad == 0 else f"FAILED on {bad} image(s)")
    return 0 if bad == 0 else 2


if __name__ == "__main__":
    sys.exit(main())
