#!/usr/bin/env python
"""The "kernel to beat" (SURVEY.md 8(d) last row): the reference's own CUDA kernels
(spotfinder/kernels/thresholding.cu, erosion.cu behind the wrappers of spotfinder/spotfinder.cu),
compiled in place for sm_100 (oracle/Makefile target `refgpu`), timed on the same device-resident
frames as this repo's kernels.  Prints one JSON object.

  wrapper_us_per_frame : call_do_spotfinding_dispersion/_extended as the reference frame loop calls
                         them (per frame: constant upload + 2-5 stream syncs [+ 2 cudaMallocPitch for
                         extended], spotfinder.cu:148-189,213-347)
  kernels_us_per_frame : the same kernels queued back to back with no host sync (device time)
  ours_us_per_frame    : ffs_threshold on a batch of the same frames (device time, CUDA events)
"""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "fast-feedback-service_b200"))

from ffs_b200 import synth  # noqa: E402
from ffs_b200.spotfinder import Spotfinder  # noqa: E402
from oracle import oracle  # noqa: E402


def ours(g, mask, frames, algorithm, batch, reps=5):
    import torch
    n = len(frames)
    with Spotfinder(g.width, g.height, frames.dtype, mask, n_slots=batch, algorithm=algorithm) as sf:
        for s0 in range(0, batch, n):
            sf.upload(frames[: min(n, batch - s0)], s0)
        st = torch.cuda.Stream()
        for _ in range(3):
            sf.threshold(0, batch, st.cuda_stream, sync=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(reps):
            sf.threshold(0, batch, st.cuda_stream, sync=False)
        e1.record(st)
        e1.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (reps * batch)


def main():
    out = {"what": "reference CUDA kernels (sm_100 build of /root/reference/spotfinder) vs this repo, same frames, "
                   "device-resident, us per frame", "cases": {}}
    cases = [
        ("eiger16m_u16_dispersion", synth.EIGER2_16M, np.uint16, "dispersion", 32),
        ("eiger16m_u32_dispersion", synth.EIGER2_16M, np.uint32, "dispersion", 16),
        ("eiger16m_u16_extended", synth.EIGER2_16M, np.uint16, "dispersion_extended", 16),
        ("eiger16m_u32_extended", synth.EIGER2_16M, np.uint32, "dispersion_extended", 16),
        ("jungfrau4m_u32_dispersion", synth.JUNGFRAU_4M, np.uint32, "dispersion", 32),
    ]
    for name, g, dtype, algo, batch in cases:
        bpp = np.dtype(dtype).itemsize
        if not oracle.refgpu_available(bpp):
            out["cases"][name] = {"error": "oracle/_ref/libffs_refgpu*.so not built"}
            continue
        mask = synth.make_mask(g)
        frames = synth.make_frames(g, mask, 4, dtype, config=2, n_hot=0)
        a = oracle.ALGO_DISPERSION if algo == "dispersion" else oracle.ALGO_DISPERSION_EXTENDED
        reps = 5
        # the reference's own defaults (min_count 3) -- what its executable runs
        _, ms_w = oracle.refgpu_run(frames, mask, a, repeats=reps, want_strong=False)
        _, ms_k = oracle.refgpu_run(frames, mask, a, repeats=reps, kernels_only=True)
        us_ours = ours(g, mask, frames, algo, batch)
        n = reps * len(frames)
        out["cases"][name] = {
            "wrapper_us_per_frame": ms_w * 1e3 / n, "kernels_us_per_frame": ms_k * 1e3 / n,
            "ours_us_per_frame": us_ours, "ours_batch": batch,
            "speedup_vs_kernels": ms_k * 1e3 / n / us_ours, "speedup_vs_wrapper": ms_w * 1e3 / n / us_ours,
        }
        print(name, out["cases"][name], file=sys.stderr)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
