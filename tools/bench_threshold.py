"""Developer timing of the threshold kernel alone (device-resident frames)."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "fast-feedback-service_b200"))
import numpy as np, torch
from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder

geom = synth.GEOMETRIES[os.environ.get("GEOM", "eiger2_16m")]
dtype = np.uint32 if os.environ.get("BPP", "2") == "4" else np.uint16
nslots = int(os.environ.get("SLOTS", "8"))
algo = os.environ.get("ALGO", "dispersion")
mask = synth.make_mask(geom)
bg = float(os.environ.get("BG", "2.0"))  # Poisson background level
frames = np.stack([synth.make_frame(geom, mask, dtype, 2, i, background=bg) for i in range(min(nslots, 4))])
frames = np.concatenate([frames] * ((nslots + len(frames) - 1) // len(frames)))[:nslots]
sf = Spotfinder(geom.width, geom.height, dtype, mask, n_slots=nslots, algorithm=algo)
sf.upload(frames)
st = torch.cuda.Stream()
fn = sf.process if os.environ.get("PROCESS") else sf.threshold
for batch in [int(b) for b in os.environ.get("BATCHES", "1,2,4,8,16,32").split(",")]:
    if batch > nslots: break
    for _ in range(3):
        fn(0, batch, st.cuda_stream, sync=True)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    reps = max(2, 64 // batch)
    with torch.cuda.stream(st):
        e0.record(st)
        for _ in range(reps):
            fn(0, batch, st.cuda_stream, sync=False)
        e1.record(st)
    e1.synchronize()
    ms = e0.elapsed_time(e1) / reps
    bthr = geom.pixels * dtype().itemsize + 2 * ((geom.pixels + 7) // 8)
    print(f"batch {batch:3d}: {ms/batch*1e3:8.1f} us/frame  {batch/ms*1e3:9.0f} fps  {bthr*batch/ms/1e6:7.0f} GB/s alg", flush=True)
