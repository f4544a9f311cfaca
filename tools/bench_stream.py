"""Developer timing of the streaming API (dense and compressed ingestion)."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "fast-feedback-service_b200"))
import numpy as np
from ffs_b200 import synth, bslz4_encode
from ffs_b200.spotfinder import Spotfinder

geom = synth.GEOMETRIES[os.environ.get("GEOM", "eiger2_16m")]
dtype = np.uint32 if os.environ.get("BPP", "2") == "4" else np.uint16
ring = int(os.environ.get("RING", "8"))
N = int(os.environ.get("N", "400"))
mask = synth.make_mask(geom)
frames = [synth.make_frame(geom, mask, dtype, 2, i) for i in range(2)]
chunks = [bslz4_encode.compress_chunk(f) for f in frames]
sf = Spotfinder(geom.width, geom.height, dtype, mask, n_slots=ring)
clen = {}

def run(n, comp, fill=False):
    t_acq = t_sub = t_wait = 0.0
    for i in range(n):
        if sf.pending >= ring:
            t = time.perf_counter(); sf.wait(16); t_wait += time.perf_counter() - t
        t = time.perf_counter()
        if comp:
            slot, buf = sf.acquire_compressed()
            if fill or slot not in clen:
                ch = chunks[slot % 2]; buf[:len(ch)] = np.frombuffer(ch, np.uint8); clen[slot] = len(ch)
            t_acq += time.perf_counter() - t; t = time.perf_counter()
            sf.submit_compressed(slot, i, clen[slot])
        else:
            slot, buf = sf.acquire()
            if fill: buf[...] = frames[slot % 2]
            t_acq += time.perf_counter() - t; t = time.perf_counter()
            sf.submit(slot, i)
        t_sub += time.perf_counter() - t
    while sf.pending:
        sf.wait(16)
    return t_acq, t_sub, t_wait

for comp in (False, True):
    run(ring * 2, comp, fill=True)
    t0 = time.perf_counter()
    ta, ts, tw = run(N, comp)
    dt = time.perf_counter() - t0
    print(f"{'compressed' if comp else 'dense':10s} ring {ring}: {N/dt:8.0f} fps  ({dt/N*1e6:6.1f} us/frame; host: acquire {ta/N*1e6:5.1f} submit {ts/N*1e6:5.1f} wait {tw/N*1e6:5.1f} us)", flush=True)

# multi-threaded producers (the reference runs N reader threads): compressed ingestion
import threading
for T in (2, 4, 8):
    per = N // T
    def producer():
        for i in range(per):
            slot, buf = sf.acquire_compressed()
            sf.submit_compressed(slot, i, clen[slot])
    run(ring * 2, True, fill=True)
    ths = [threading.Thread(target=producer) for _ in range(T)]
    t0 = time.perf_counter()
    for t in ths: t.start()
    got = 0
    while got < per * T:
        if sf.pending:
            sf.wait(16); got += 1
        else:
            time.sleep(0.00002)
    dt = time.perf_counter() - t0
    for t in ths: t.join()
    print(f"compressed, {T} producer threads: {per*T/dt:8.0f} fps ({dt/(per*T)*1e6:6.1f} us/frame)", flush=True)
