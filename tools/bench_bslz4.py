"""Developer timing of the device bitshuffle-LZ4 decode kernel alone."""
import os, sys, time
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
out = sf.decode_chunk(chunk)
assert np.array_equal(out, f)
L = sf._L
for _ in range(3):
    L.ffs_decode_chunk_device(sf._ctx, 0, chunk, len(chunk), None)
torch.cuda.synchronize()
t0 = time.perf_counter()
N = 50
for _ in range(N):
    L.ffs_decode_chunk_device(sf._ctx, 0, chunk, len(chunk), None)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / N
print(f"decode_chunk_device (memcpy+index+H2D+kernel+sync): {dt*1e6:.1f} us/frame, chunk {len(chunk)/1e6:.2f} MB ratio {f.nbytes/len(chunk):.2f}")
