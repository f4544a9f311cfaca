"""Developer tool: per-warp timeline of the device bitshuffle-LZ4 decoder (library built with
-DBSLZ4_TIMELINE, see csrc/bslz4_gpu.cuh): when each block started / finished its LZ4 stage /
finished, on which SM -- shows how much of the kernel is the tail of a partly filled wave."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "fast-feedback-service_b200"))
import numpy as np, torch, ctypes as C
from ffs_b200 import synth, bslz4_encode
from ffs_b200.spotfinder import Spotfinder

geom = synth.GEOMETRIES[os.environ.get("GEOM", "eiger2_16m")]
dtype = np.uint32 if os.environ.get("BPP", "2") == "4" else np.uint16
mask = synth.make_mask(geom)
f = synth.make_frame(geom, mask, dtype, 2, 0)
chunk = bslz4_encode.compress_chunk(f)
sf = Spotfinder(geom.width, geom.height, dtype, mask, n_slots=2)
L = sf._L
for _ in range(3):
    L.ffs_decode_chunk_device(sf._ctx, 0, chunk, len(chunk), None)
torch.cuda.synchronize()
tl = np.zeros((16384, 4), np.uint32)
L.ffs_debug_bslz4_timeline.argtypes = [C.c_void_p, C.c_size_t]
rc = L.ffs_debug_bslz4_timeline(tl.ctypes.data, tl.nbytes)
assert rc == 0, rc
tl = tl[tl[:, 2] != 0]
t0 = tl[:, 0].min()
start = (tl[:, 0] - t0).astype(np.int64); lz = (tl[:, 1] - t0).astype(np.int64); end = (tl[:, 2] - t0).astype(np.int64)
smid = tl[:, 3] >> 16
print(f"warps {len(tl)}  kernel span {end.max()/1e3:.1f} us")
print("start pct (us):", np.percentile(start, [0, 25, 50, 75, 87, 90, 95, 99, 100]) / 1e3)
print("end   pct (us):", np.percentile(end, [0, 25, 50, 75, 90, 95, 99, 100]) / 1e3)
dur = end - start
first = start < np.percentile(start, 50)
late = start > np.percentile(start, 90)
print(f"duration us: all {dur.mean()/1e3:.1f}  started-early {dur[first].mean()/1e3:.1f}  started-late {dur[late].mean()/1e3:.1f}  max {dur.max()/1e3:.1f}")
print(f"lz4 share of a block: {((lz-start).sum()/dur.sum()):.2f}")
per_sm_end = np.array([end[smid == s].max() for s in np.unique(smid)])
print("per-SM last end (us) pct:", np.percentile(per_sm_end, [0, 10, 50, 90, 100]) / 1e3, " SMs", len(per_sm_end))
# resident warps over time
ts = np.arange(0, end.max(), 2000)
res = [(int(((start <= t) & (end > t)).sum())) for t in ts]
print("resident warps every 2 us:", res)
