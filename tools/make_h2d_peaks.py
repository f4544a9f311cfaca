#!/usr/bin/env python
"""Condense tools/ubench/h2d outputs (one JSON per GPU count) into profiles/h2d_peaks.json, the
ceiling bench.py reports `e2e.h2d_frac` against.

    python tools/make_h2d_peaks.py gpurun_out/r2n8/h2d_n1.json gpurun_out/r2n8/h2d_n2.json ...

per_gpu_ceiling_gbs[N] = the best pure-copy rate per GPU of a frame-sized contiguous
cudaMemcpyAsync (what ffs_submit issues) with N GPUs copying at once -- the slowest GPU's rate,
since the stream is as fast as its slowest shard."""
import json
import sys
from pathlib import Path

out = {"source": "tools/ubench/h2d.cu (pinned host -> device, CUDA events per GPU, all GPUs started together)",
       "per_gpu_ceiling_gbs": {}, "aggregate_ceiling_gbs": {}, "runs": {}}
for f in sys.argv[1:]:
    d = json.loads(Path(f).read_text())
    n = str(d["gpus"])
    v = d["variants"]
    best = max(("1d_frame", "1d_frame_2s"), key=lambda k: min(v[k]["per_gpu_gbs"]))
    out["per_gpu_ceiling_gbs"][n] = round(min(v[best]["per_gpu_gbs"]), 2)
    out["aggregate_ceiling_gbs"][n] = round(v[best]["aggregate_gbs"], 2)
    out["runs"][n] = {k: {"per_gpu_min": round(min(x["per_gpu_gbs"]), 2), "per_gpu_max": round(max(x["per_gpu_gbs"]), 2),
                          "aggregate": x["aggregate_gbs"], **({"host_writer_gbs": x["host_writer_gbs"]}
                                                               if "host_writer_gbs" in x else {})}
                      for k, x in v.items()}
    out["frame_bytes"] = d["frame_bytes"]
Path(__file__).resolve().parents[1].joinpath("profiles", "h2d_peaks.json").write_text(json.dumps(out, indent=1) + "\n")
print(json.dumps(out["per_gpu_ceiling_gbs"]), json.dumps(out["aggregate_ceiling_gbs"]))
