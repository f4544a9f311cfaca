"""Field parity check for a stream directory or a NeXus file -- the role of the reference's `--validate`
(spotfinder.cc:1012-1052: re-run the CPU algorithm on each image and compare the strong pixels).

The GPU executable contains no CPU implementation (the product path must fail rather than fall
back), so the comparison lives here, in the test tree: the unmodified `spotfinder` executable
is run with --writeout on the directory, and every pixels_NNNNN.txt it writes is compared with
the CPU oracle's DIALS dispersion of the same decoded frame.

    python tests/validate_stream.py <stream dir | FILE.nxs> [--images N] [--algorithm dispersion|dispersion_extended]
                                    [-- extra arguments for the executable, e.g. --detector JSON]

A NeXus file is read here through the same in-tree reader the executable uses (`h5lite_tool h5read`:
metadata, mask, the stored chunk of each frame), decoded by the oracle's bitshuffle-LZ4 port.

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


class StreamSource:
    def __init__(self, d: Path):
        self.d = d
        hdr = json.loads((d / "start_1").read_text())
        self.h, self.w = hdr["y_pixels_in_detector"], hdr["x_pixels_in_detector"]
        self.bits = hdr["bit_depth_image"]
        self.dtype = np.uint16 if self.bits == 16 else np.uint32
        self.mask = (np.frombuffer((d / "start_5").read_bytes(), np.int32).reshape(self.h, self.w) == 0).astype(np.uint8)
        self.n = hdr["nimages"]
        # trusted range: pixels above the (pixel_t-narrowed) count cutoff are invalid
        self.cutoff = int(np.array(int(hdr.get("countrate_correction_count_cutoff", 0))).astype(self.dtype))

    def frame(self, i):
        return decompress_chunk((self.d / f"image_{i:06d}_2").read_bytes(), self.dtype, self.h * self.w).reshape(self.h, self.w)


class NexusSource:
    TOOL = ROOT / "fast-feedback-service_b200" / "ffs_b200" / "h5lite_tool"

    def __init__(self, path: Path):
        self.path = path
        r = subprocess.run([str(self.TOOL), "h5read", str(path)], capture_output=True, check=True)
        info = json.loads(r.stdout.decode().strip().splitlines()[-1])
        self.h, self.w = info["shape"]
        self.bits = 8 * info["element_size"]
        self.dtype = np.uint16 if self.bits == 16 else np.uint32
        self.n = info["frames"]
        self.raw = info["raw_pixels"]
        self.cutoff = int(np.array(info["trusted"][1]).astype(self.dtype))
        m = subprocess.run([str(self.TOOL), "cat", str(path), "/entry/instrument/detector/pixel_mask"], capture_output=True)
        if m.returncode == 0 and len(m.stdout) in (4 * self.h * self.w, 8 * self.h * self.w):
            raw = np.frombuffer(m.stdout, np.uint32 if len(m.stdout) == 4 * self.h * self.w else np.uint64)
            self.mask = (raw.reshape(self.h, self.w) == 0).astype(np.uint8)
        else:
            self.mask = np.ones((self.h, self.w), np.uint8)

    def frame(self, i):
        r = subprocess.run([str(self.TOOL), "h5read", str(self.path), str(i)], capture_output=True, check=True)
        if self.raw:
            return np.frombuffer(r.stdout, self.dtype).reshape(self.h, self.w)
        return oracle.bslz4_decode(r.stdout, self.dtype, self.h * self.w).reshape(self.h, self.w)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("stream")
    ap.add_argument("--images", type=int, default=0)
    ap.add_argument("--algorithm", default="dispersion")
    ap.add_argument("extra", nargs="*", help="after --: passed to the executable")
    a = ap.parse_args()
    d = Path(a.stream)
    src = StreamSource(d) if d.is_dir() else NexusSource(d)
    h, w, bits, dtype, mask = src.h, src.w, src.bits, src.dtype, src.mask
    n = a.images or src.n
    exe = ROOT / "fast-feedback-service_b200" / "ffs_b200" / ("spotfinder" if bits == 16 else "spotfinder32")
    with tempfile.TemporaryDirectory() as tmp:
        cmd = [str(exe), str(d.resolve()), "--images", str(n), "--threads", "2", "--writeout", "--algorithm", a.algorithm,
               *a.extra]
        r = subprocess.run(cmd, cwd=tmp, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stdout[-2000:], r.stderr[-2000:])
            return r.returncode or 1
        bad = 0
        for i in range(n):
            frame = src.frame(i)
            fn = oracle.dispersion if a.algorithm == "dispersion" else oracle.dispersion_extended
            cutoff = src.cutoff
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
