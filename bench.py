#!/usr/bin/env python
"""bench.py -- frames/s of the DIALS dispersion spot-finding hot path on B200.

Default workload = BASELINE.json configs[1]: Eiger2 16M (4148x4362) 16-bit frames, standard
dispersion (7x7, nsig_b 6, nsig_s 3, min_count 2) + 2D connected components + per-spot
reduction, synthetic Poisson background + Gaussian spots + module-gap mask, frames sharded
frame-wise across ranks (no data-path collective: "scaling": "weak").

  value : whole-job frames/s with the frames already resident in HBM (slot API:
          ffs_process on a batch of B frames per step; B frames are far larger than L2).
  e2e   : the same metric through the streaming C-ABI calls (ffs_acquire / ffs_submit /
          ffs_wait) with HOST buffers: every frame is copied from pinned host memory to the
          device inside the timed region and its counters + spot list are read back.
          `h2d_frac` = achieved H2D GB/s per GPU over the measured pure-copy ceiling for this
          number of concurrently active GPUs (tools/ubench/h2d.cu -> profiles/h2d_peaks.json).
  e2e_from_chunks : the streaming path fed with the raw bitshuffle-LZ4 chunk (what the reference's
          reader returns, spotfinder.cc:807-834), decoded on the device.  The reference arm has the
          same leg (host decode + threshold + CC).
  roofline    : the threshold kernel(s), bracketed by CUDA events on their own stream inside the
                timed region; algorithmic bytes per frame = image + 1-bit validity plane + 1-bit
                strong plane.
  cpu_baseline: the reference's CPU path on this box's host cores (rank 0, N=1, bounded sample).
  reference_gpu: the reference's own CUDA kernels (sm_100 build, oracle/_ref/libffs_refgpu*.so) on
                the same device-resident frames: the kernel to beat.
  extra       : the other BASELINE.json configurations, one bounded leg each (cfg 3 extended u32 with
                its own roofline, cfg 5 Jungfrau 4M stream rate + paced 2500 Hz latency p50/p99,
                cfg 4 rotation sweep at N > 1).

`--config NAME` makes another configuration the headline line (same keys):
  eiger16m_u16 (default) | eiger4m_u16 (cfg 1) | eiger16m_u32_ext (cfg 3) | sweep3600 (cfg 4) |
  jungfrau4m_u32 (cfg 5).
`--impl reference` times the reference's CPU implementation of the chosen configuration with all
host threads (rank 0 only).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "fast-feedback-service_b200"))

from ffs_b200 import synth  # noqa: E402

HEADLINE_METRIC = "frames/sec (Eiger2 16M 16-bit)"  # BASELINE.json

CONFIGS = {
    # name: geometry, pixel type, algorithm, frames per step, distinct synthetic frames, description
    "eiger16m_u16": dict(geom=synth.EIGER2_16M, dtype=np.uint16, algorithm="dispersion", batch=32, distinct=8,
                         metric=HEADLINE_METRIC, cfg=2,
                         what="Eiger2 16M (4148x4362) u16: dispersion 7x7 nsig_b=6 nsig_s=3 min_count=2 + 2D connected "
                              "components + spot reduction"),
    "eiger4m_u16": dict(geom=synth.EIGER2_4M, dtype=np.uint16, algorithm="dispersion", batch=64, distinct=8,
                        metric="frames/sec (Eiger2 4M 2070x2167 16-bit)", cfg=1,
                        what="Eiger2 4M (2070x2167) u16: dispersion 7x7 nsig_b=6 nsig_s=3 min_count=2 + 2D connected "
                             "components (BASELINE.json configs[0], the reference's CPU-runnable case)"),
    "eiger16m_u32_ext": dict(geom=synth.EIGER2_16M, dtype=np.uint32, algorithm="dispersion_extended", batch=16,
                             distinct=4, metric="frames/sec (Eiger2 16M 32-bit, dispersion_extended)", cfg=3,
                             what="Eiger2 16M (4148x4362) u32 (PIXEL_DATA_32BIT): dispersion_extended two-pass threshold "
                                  "with module-gap mask + 2D connected components"),
    "jungfrau4m_u32": dict(geom=synth.JUNGFRAU_4M, dtype=np.uint32, algorithm="dispersion", batch=32, distinct=8,
                           metric="frames/sec (Jungfrau 4M 32-bit)", cfg=5,
                           what="Jungfrau 4M (2068x2164) u32: dispersion + 2D connected components, per-frame spot "
                                "list + centroids (small-frame latency path, 2500 Hz target)"),
    "sweep3600": dict(geom=synth.EIGER2_16M, dtype=np.uint16, algorithm="dispersion", batch=32, distinct=8,
                      metric="frames/sec (Eiger2 16M 16-bit rotation sweep, 3D connected components)", cfg=4,
                      what="3D connected components over a 3600-frame synthetic rotation sweep (Eiger2 16M u16), "
                           "contiguous frame slabs per GPU, boundary-plane label merge"),
}


def algorithmic_bytes_per_frame(g, dtype) -> int:
    """image read once + 1-bit validity plane read + 1-bit strong plane written."""
    return g.pixels * np.dtype(dtype).itemsize + 2 * ((g.pixels + 7) // 8)


def make_inputs(cfg, n_distinct=None):
    g = cfg["geom"]
    n = n_distinct or cfg["distinct"]
    mask = synth.make_mask(g)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        frames = list(ex.map(lambda i: synth.make_frame(g, mask, cfg["dtype"], cfg["cfg"], i), range(n)))
    return mask, np.stack(frames)


def workload_config(cfg, name, batch, n_distinct):
    return {"workload": cfg["what"] + "; synthetic Poisson(2)+Gaussian spots, module-gap mask; frame-wise sharding",
            "name": name, "frames_per_step": batch,
            "l2": "inputs larger than L2 (a step's batch of frames >> 126 MB)",
            "distinct_frames": n_distinct}


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU implementation of the path
# ---------------------------------------------------------------------------------------------
class CpuPath:
    """The reference's own CPU code for every step it has some, compiled in place under oracle/_ref
    (oracle/Makefile): threshold = standalone.cc (standard dispersion) or baseline.cpp's
    DispersionExtendedThreshold (dispersion_extended); connected components + both filters =
    connected_components.cc (ConnectedComponents + find_2d_components, one object per frame as
    spotfinder.cc:897-926 builds them).  Where a library is not built the oracle port stands in
    and `kind` says "port".  Chunk decode (from_chunks): the oracle's scalar bitshuffle-LZ4 port
    (the reference links the bitshuffle library, which is not in the image)."""

    def __init__(self, mask, n_threads, algorithm="dispersion"):
        from oracle import oracle
        self.o = oracle
        self.mask = np.ascontiguousarray(mask, np.uint8)
        self.algorithm = algorithm
        self.n_threads = n_threads
        self.h, self.w = mask.shape
        self.local = threading.local()
        self.L = oracle.lib()
        self.R = oracle.ref() if (algorithm == "dispersion" and oracle.ref_available()) else None
        self.X = oracle.refext() if (algorithm != "dispersion" and oracle.refext_available()) else None
        self.cc_ok = oracle.refcc_available(2) and oracle.refcc_available(4)
        self.kind = "reference" if ((self.R or self.X) and self.cc_ok) else "port"
        # ONE pool for the object's lifetime: a reference context is ~0.6 GB at 16M and lives in
        # its worker thread, so a pool per run() would leak n_threads contexts per step.
        self.ex = ThreadPoolExecutor(max_workers=n_threads)
        self.ctxs = []
        self.lock = threading.Lock()

    def describe(self):
        thr = ("reference standalone.cc compiled in place (oracle/_ref)" if self.R
               else "reference baseline.cpp DispersionExtendedThreshold compiled in place (oracle/_ref)" if self.X
               else "oracle port")
        cc = ("reference connected_components.cc compiled in place (oracle/_ref; stand-in Boost.Graph)" if self.cc_ok
              else "oracle port")
        return f"threshold = {thr}, CC + filters = {cc}"

    def _ctx(self):
        if not hasattr(self.local, "ctx"):
            self.local.dst = np.empty((self.h, self.w), np.uint8)
            self.local.spots = np.zeros(1 << 16, self.o.REFCC_SPOT_DTYPE)
            if self.R:
                self.local.ctx = self.R.ffs_ref_create(self.w, self.h)
            elif self.X:
                p = self.o.default_params()
                self.local.ctx = self.X.ffs_refext_create(1, self.w, self.h, self.mask.ctypes.data, p.kernel_half_x,
                                                          p.kernel_half_y, p.nsig_b, p.nsig_s, p.threshold, p.min_count)
            else:
                self.local.ctx = None
            if self.local.ctx:
                with self.lock:
                    self.ctxs.append(self.local.ctx)
        return self.local.ctx, self.local.dst

    def close(self):
        self.ex.shutdown(wait=True)
        for c in self.ctxs:
            if self.R:
                self.R.ffs_ref_free(c)
            else:
                self.X.ffs_refext_destroy(c)
        self.ctxs = []

    def one(self, frame):
        ctx, dst = self._ctx()
        if isinstance(frame, tuple):  # (chunk bytes, dtype): decode first, as spotfinder.cc:823-841
            chunk, dtype = frame
            frame = self.o.bslz4_decode(chunk, dtype, self.h * self.w).reshape(self.h, self.w)
        if self.R:
            self.R.ffs_ref_standard_dispersion(ctx, frame.ctypes.data, frame.dtype.itemsize,
                                               self.mask.ctypes.data, dst.ctypes.data)
        elif self.X:
            self.X.ffs_refext_run(ctx, frame.ctypes.data, frame.dtype.itemsize, dst.ctypes.data)
        else:
            p = self.o.default_params()
            fn = (self.L.ffs_oracle_dispersion if self.algorithm == "dispersion"
                  else self.L.ffs_oracle_dispersion_extended)
            fn(frame.ctypes.data, frame.dtype.itemsize, self.mask.ctypes.data, self.w, self.h, C.byref(p),
               dst.ctypes.data)
        if self.cc_ok:
            G = self.o.refcc(frame.dtype.itemsize)
            h = G.ffs_refcc_new()
            try:
                G.ffs_refcc_add_slice(h, dst.ctypes.data, frame.ctypes.data, self.w, self.h, 3, None, None, None)
                n = G.ffs_refcc_find_2d(h, 0, 3, 2.0, self.local.spots.ctypes.data, len(self.local.spots), None)
            finally:
                G.ffs_refcc_free(h)
            return int(n)
        _, spots = self.o.connected_components(dst, frame, row_wrap=True, want_labels=False)
        kept, _, _ = self.o.filter_spots(spots, 3, 2.0)
        return len(kept)

    def run(self, frames, n_frames):
        """Process n_frames (cycled) frame-parallel over n_threads; returns seconds."""
        ex = self.ex
        if len(self.ctxs) < self.n_threads and (self.R or self.X):  # first call: let every worker build its context
            list(ex.map(self.one, [frames[i % len(frames)] for i in range(self.n_threads)]))
        t0 = time.perf_counter()
        list(ex.map(self.one, [frames[i % len(frames)] for i in range(n_frames)]))
        return time.perf_counter() - t0


def cpu_threads(g) -> int:
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    try:  # 24-byte SAT cell + double image + u8 planes per pixel and context
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
        n = min(n, max(1, int(avail * 0.5 / (40.0 * g.pixels))))
    except Exception:
        pass
    return max(1, min(n, 64))


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    g = cfg["geom"]
    n_distinct = cfg["distinct"]  # the same frames, and the same frames per step, as the GPU arm's config
    mask, frames = make_inputs(cfg, n_distinct)
    T = cpu_threads(g)
    cpu = CpuPath(mask, T, cfg["algorithm"])
    per_step = args.batch or cfg["batch"]  # a bounded sample of the stream, frame-parallel over T threads
    for _ in range(args.warmup):
        cpu.run(frames, per_step)
    t = sum(cpu.run(frames, per_step) for _ in range(args.steps))
    fps = per_step * args.steps / t
    sample = f"{per_step} frames/step x {args.steps} steps, frame-parallel over {T} threads; {cpu.describe()}"
    out = {
        "impl": "reference", "metric": cfg["metric"], "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(cfg, args.config, per_step, n_distinct),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": T, "kind": cpu.kind, "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if args.config == "sweep3600":  # + 3D connected components of the sampled frames, single thread
        from oracle import oracle
        strong = np.stack([oracle.dispersion(f, mask) for f in frames])
        if oracle.refcc_available(frames.dtype.itemsize):  # the reference's own find_3d_components
            with oracle.RefCC(frames, strong, 3) as r:
                t0 = time.perf_counter()
                r.find_3d(3, 2.0)
                t3 = (time.perf_counter() - t0) / len(frames)
            out["cpu_baseline"]["cc3d_code"] = "reference find_3d_components (oracle/_ref/libffs_refcc)"
        else:
            t0 = time.perf_counter()
            oracle.connected_components(strong, frames, want_labels=False)
            t3 = (time.perf_counter() - t0) / len(frames)
            out["cpu_baseline"]["cc3d_code"] = "oracle port"
        out["cpu_baseline"]["cc3d_s_per_frame_single_thread"] = t3
        out["value"] = out["e2e"]["value"] = out["cpu_baseline"]["value"] = 1.0 / (1.0 / fps + t3)
    if not args.no_chunks:  # the real loop starts from the compressed chunk (spotfinder.cc:807-834)
        from ffs_b200 import bslz4_encode
        with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
            chunks = list(ex.map(bslz4_encode.compress_chunk, frames))
        items = [(c, frames.dtype) for c in chunks]
        cpu.run(items, min(T, 4))
        tc = cpu.run(items, per_step)
        out["e2e_from_chunks"] = {"value": per_step / tc, "unit": "frames/s",
                                  "what": "host bitshuffle-LZ4 decode (oracle scalar port of bshuf_decompress_lz4) + "
                                          "threshold + CC, frame-parallel", "cores": T}
    cpu.close()
    emit(out)


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.idx = device_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out = self.proc.communicate(timeout=5)[0]
        except Exception:
            self.proc.kill()
            out = ""
        mhz, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                mhz.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(mhz)) if mhz else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(mhz)}


def measured_peak_hbm():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def h2d_ceiling(world):
    """Pure-copy H2D GB/s per GPU with `world` GPUs copying at once (tools/ubench/h2d.cu, the 2D
    pitched frame copy the streaming path issues); None until it has been measured."""
    try:
        d = json.loads((ROOT / "profiles" / "h2d_peaks.json").read_text())
        return float(d["per_gpu_ceiling_gbs"][str(world)]), d.get("source", "profiles/h2d_peaks.json")
    except Exception:
        return None, None


def ncu_traffic_per_frame(key):
    """dram bytes (read+write) per frame of the threshold kernel(s) from the committed ncu capture."""
    try:
        return float(json.loads((ROOT / "profiles" / "roofline_traffic.json").read_text())[key])
    except Exception:
        return None


def parity_check(sf, cfg, mask, frame):
    """One frame against the CPU oracle before any number is emitted (slot 0 holds `frame`)."""
    from oracle import oracle
    fn = oracle.dispersion if cfg["algorithm"] == "dispersion" else oracle.dispersion_extended
    want = fn(frame, mask)
    sf.process(0, 1)
    got = sf.strong_mask(0)
    res, spots = sf.fetch(0)
    _, spots_want = oracle.connected_components(want, frame, want_labels=False)
    ok = bool(np.array_equal(got, want) and res.n_components == len(spots_want)
              and np.array_equal(spots["label"], spots_want["label"])
              and np.array_equal(spots["num_pixels"], spots_want["num_pixels"])
              and np.allclose(spots["com_x"], spots_want["com_x"], rtol=1e-5))
    if not ok:
        raise SystemExit("bench.py: the GPU result differs from the CPU oracle on the check frame -- no number emitted")
    return {"frame": 0, "strong_pixels": int(want.sum()), "components": int(len(spots_want)),
            "identical_to_oracle": True}


def stream_dense(sf, n, ring, frames=None, fill=False):
    """n frames through acquire -> (fill) -> submit -> wait; at most `ring` in flight."""
    for i in range(n):
        if sf.pending >= ring:
            sf.wait(16)
        slot, buf = sf.acquire()
        if fill:  # only in the warm-up: afterwards every pinned ring slot holds a frame
            buf[...] = frames[i % len(frames)]
        sf.submit(slot, i)
    last = None
    while sf.pending:
        last, _ = sf.wait(16)
    return last


def stream_with_producers(sf, frames, n, n_prod):
    """The real loop has N host threads WRITING each frame into the pinned ring (the decompressor's
    output, spotfinder.cc:828-834); here n_prod threads copy a dense frame into the slot they
    acquired and submit, this thread collects.  Returns seconds."""
    per = n // n_prod

    def producer(k):
        for i in range(per):
            slot, buf = sf.acquire()
            np.copyto(buf, frames[(k + i) % len(frames)])
            sf.submit(slot, k + i * n_prod)

    threads = [threading.Thread(target=producer, args=(k,)) for k in range(n_prod)]
    t0 = time.perf_counter()
    for t in threads:
        t.start()
    got = 0
    while got < per * n_prod:
        if sf.pending:
            sf.wait(16)
            got += 1
        else:
            time.sleep(2e-5)
    dt = time.perf_counter() - t0
    for t in threads:
        t.join()
    return dt, per * n_prod


def stream_chunks_leg(sf, frames, ring, n_total, n_prod=4):
    """Compressed ingestion: raw bitshuffle-LZ4 chunks cross PCIe and are decoded on the device."""
    from ffs_b200 import bslz4_encode
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        chunks = list(ex.map(bslz4_encode.compress_chunk, frames))
    chunk_len = {}

    def serial(n, fill=False):
        for i in range(n):
            if sf.pending >= ring:
                sf.wait(16)
            slot, buf = sf.acquire_compressed()
            if fill or slot not in chunk_len:
                ch = chunks[slot % len(chunks)]
                buf[:len(ch)] = np.frombuffer(ch, np.uint8)
                chunk_len[slot] = len(ch)
            sf.submit_compressed(slot, i, chunk_len[slot])
        last = None
        while sf.pending:
            last, _ = sf.wait(16)
        return last

    serial(ring * 2, fill=True)  # warm-up; every ring slot now holds a chunk
    # Producer threads as in the reference's frame loop (spotfinder.cc:725: N reader threads share
    # the ring): each indexes its chunk's block headers and submits; this thread collects results.
    n = (n_total // n_prod) * n_prod

    def producer(k):
        for i in range(n // n_prod):
            slot, _ = sf.acquire_compressed()
            sf.submit_compressed(slot, k + i * n_prod, chunk_len[slot])

    threads = [threading.Thread(target=producer, args=(k,)) for k in range(n_prod)]
    t0 = time.perf_counter()
    for t in threads:
        t.start()
    got, last = 0, None
    while got < n:
        if sf.pending:
            last, _ = sf.wait(16)
            got += 1
        else:
            time.sleep(2e-5)
    dt = time.perf_counter() - t0
    for t in threads:
        t.join()
    assert last.status == 0 and last.num_strong_pixels > 0
    return dt, n, float(np.mean([len(c) for c in chunks]))


def measure_config(cfg, name, args, env, full=True):
    """value / e2e / roofline of one frame-stream configuration on this rank; returns a dict of
    rank-local numbers (times in ms) plus static facts.  full=False: the bounded `extra` variant."""
    import torch
    from ffs_b200.spotfinder import Spotfinder
    g, dtype, B = cfg["geom"], cfg["dtype"], (args.batch or cfg["batch"])
    steps, warmup = (args.steps, max(args.warmup, 3)) if full else (max(3, args.steps // 4), 3)
    mask, frames = make_inputs(cfg) if full else make_inputs(cfg, min(4, cfg["distinct"]))
    sf = Spotfinder(g.width, g.height, dtype, mask, n_slots=B, device=env["local"], algorithm=cfg["algorithm"])
    for s0 in range(0, B, len(frames)):
        sf.upload(frames[: min(len(frames), B - s0)], s0)
    r = {"B": B, "steps": steps, "warmup": warmup, "n_distinct": len(frames)}
    if env["rank"] == 0:
        r["parity_check"] = parity_check(sf, cfg, mask, frames[0])
    stream = torch.cuda.Stream()
    sh = stream.cuda_stream
    for _ in range(warmup):
        sf.process(0, B, sh, sync=True)
    sampler = ClockSampler(env["local"])
    env["barrier"]()
    if env["rank"] == 0 and full:
        sampler.start()
    sf.profile_enable(True)
    l0 = sf.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        sf.process(0, B, sh, sync=False)
    e1.record(stream)
    e1.synchronize()
    env["barrier"]()
    r["launches"] = sf.kernel_launches - l0
    r["ms"] = e0.elapsed_time(e1)
    k_ms, k_n = sf.profile_read()
    r["k_avg"] = k_ms / max(k_n, 1)
    sf.profile_enable(False)
    if env["rank"] == 0 and full:
        r["clocks"] = sampler.stop()
    res, _ = sf.fetch(0, 16)
    assert res.status == 0 and res.num_strong_pixels > 0
    sf.close()

    # ---- e2e: streaming API, pinned host frames -> H2D -> kernels -> results on host ----------
    ring = min(B, 8)
    sf = Spotfinder(g.width, g.height, dtype, mask, n_slots=ring, device=env["local"], algorithm=cfg["algorithm"])
    n_e2e = B * steps
    stream_dense(sf, ring * 2, ring, frames, fill=True)  # warm-up, and fills the pinned ring
    env["barrier"]()
    t0 = time.perf_counter()
    last = stream_dense(sf, n_e2e, ring)
    torch.cuda.synchronize()
    r["t_e2e"] = time.perf_counter() - t0
    r["n_e2e"] = n_e2e
    assert last.status == 0 and last.num_strong_pixels > 0
    env["barrier"]()
    if full:
        # producers writing the ring inside the timed region
        n_prod = min(8, max(2, (os.cpu_count() or 2) // max(1, env["world"]) // 2))
        stream_with_producers(sf, frames, ring * 2, 2)
        env["barrier"]()
        r["t_prod"], r["n_prod_frames"] = stream_with_producers(sf, frames, n_e2e, n_prod)
        r["n_prod"] = n_prod
        env["barrier"]()
        if not args.no_chunks:
            r["t_e2c"], r["n_e2c"], r["mean_chunk"] = stream_chunks_leg(sf, frames, ring, n_e2e * 4)
            env["barrier"]()
    sf.close()
    r["mask"], r["frames"] = mask, frames
    return r


def latency_leg(cfg, mask, frames, device, rate=2500.0, count=5000):
    """cfg 5: paced-input per-frame latency through the C ABI (native probe, csrc/cli/latency_main.cc)."""
    exe = ROOT / "fast-feedback-service_b200" / "ffs_b200" / "ffs_latency"
    if not exe.exists():
        return {"error": "ffs_latency not built"}
    g = cfg["geom"]
    with tempfile.TemporaryDirectory() as tmp:
        fp, mp = Path(tmp) / "frames.raw", Path(tmp) / "mask.raw"
        frames.tofile(fp)
        np.ascontiguousarray(mask, np.uint8).tofile(mp)
        cmd = [str(exe), "--frames", str(fp), "--mask", str(mp), "--width", str(g.width), "--height", str(g.height),
               "--bpp", str(frames.dtype.itemsize), "--n-distinct", str(len(frames)), "--count", str(count),
               "--rate", str(rate), "--ring", "8", "--device", str(device),
               "--algorithm", "0" if cfg["algorithm"] == "dispersion" else "1"]
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    if p.returncode != 0:
        return {"error": (p.stderr or p.stdout)[-300:]}
    out = json.loads(p.stdout.strip().splitlines()[-1])
    out["what"] = ("submit -> result-on-host latency per frame at a paced input rate, streaming C ABI, frames in "
                   "the pinned ring (measured from each frame's scheduled time)")
    return out


def reference_gpu_leg(cases):
    """The reference's own CUDA kernels on the same frames (oracle/_ref/libffs_refgpu*.so)."""
    from oracle import oracle
    out = {"what": "reference spotfinder/kernels/*.cu compiled in place for sm_100, device-resident frames, us per "
                   "frame by CUDA events; `wrapper` = call_do_spotfinding_* as the frame loop calls them (host "
                   "syncs, constant upload [, 2 cudaMallocPitch]), `kernels` = the same kernels queued back to back"}
    for key, (frames, mask, algo, ours_us) in cases.items():
        bpp = frames.dtype.itemsize
        if not oracle.refgpu_available(bpp):
            out[key] = {"unavailable": "oracle/_ref/libffs_refgpu*.so not built"}
            continue
        try:
            a = oracle.ALGO_DISPERSION if algo == "dispersion" else oracle.ALGO_DISPERSION_EXTENDED
            reps = 3
            _, ms_w = oracle.refgpu_run(frames, mask, a, repeats=reps, want_strong=False)
            _, ms_k = oracle.refgpu_run(frames, mask, a, repeats=reps, kernels_only=True)
            n = reps * len(frames)
            out[key] = {"wrapper_us_per_frame": ms_w * 1e3 / n, "kernels_us_per_frame": ms_k * 1e3 / n,
                        "ours_us_per_frame": ours_us, "speedup_vs_kernels": ms_k * 1e3 / n / ours_us}
        except Exception as e:  # the checker must never take the bench down
            out[key] = {"error": str(e)[:200]}
    return out


def sweep_leg(cfg, env, n_frames_total):
    """cfg 4: every rank streams its contiguous slab with sweep recording on, labels it on its GPU
    and the slabs are merged across ranks.  Returns rank-local phase times."""
    import torch
    sys.path.insert(0, str(ROOT / "tools"))
    from bench_sweep import SweepSource, run_slab
    from ffs_b200.spotfinder import Spotfinder
    from ffs_b200.sweep import NcclComm, merge_native, merge_slabs_native, slab_bounds
    g = cfg["geom"]
    mask = synth.make_mask(g)
    src = SweepSource(g, mask, n_frames_total, cfg["dtype"])
    z0, z1 = slab_bounds(n_frames_total, env["world"], env["rank"])
    times = {}
    comm = NcclComm.from_torch(env["local"]) if env["world"] > 1 else None
    with Spotfinder(g.width, g.height, cfg["dtype"], mask, n_slots=6, device=env["local"]) as sf:
        slab = run_slab(sf, src, z0, z1, times)
        env["barrier"]()
        tm = time.perf_counter()
        if comm is not None:  # native merge over NCCL (csrc/api_multi.inl)
            merged, times["exchange_s"] = merge_native(sf, comm, z0, z1 - z0, slab.spots, gather=True)
        else:
            merged, times["exchange_s"] = merge_slabs_native([slab], g.width, g.height), 0.0
        torch.cuda.synchronize()
        times["merge_s"] = time.perf_counter() - tm
        kept, _, _ = sf.filter_spots(merged, 3, 2.0)
    times.update(frames=n_frames_total, components=int(len(merged)), spots_after_filters=int(len(kept)),
                 slab_components=int(len(slab.spots)))
    if comm is not None:
        comm.close()
    return times


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(vals):
        if world == 1:
            return [float(v) for v in vals]
        t = torch.tensor(vals, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t]

    env = {"rank": rank, "world": world, "local": local, "barrier": barrier}
    name = args.config
    cfg = CONFIGS[name]
    g, dtype = cfg["geom"], cfg["dtype"]
    frame_bytes = g.pixels * np.dtype(dtype).itemsize
    peak, peak_src = measured_peak_hbm()
    out = None

    if name == "sweep3600":
        t = sweep_leg(cfg, env, args.sweep_frames)
        stream_s, synth_s, finish_s, merge_s, exch_s = reduce_max([t["stream_s"], t["synth_s"], t["finish_s"],
                                                                   t["merge_s"], t.get("exchange_s", 0.0)])
        if rank == 0:
            total = stream_s + finish_s + merge_s
            out = {"metric": cfg["metric"], "value": t["frames"] / total, "unit": "frames/s", "n_gpus": world,
                   "steps": 1, "warmup": 0, "ms_per_step": total * 1e3, "higher_is_better": True, "scaling": "strong",
                   "vs_baseline": None, "dtype": "u16", "data": "synthetic",
                   "config": {"workload": cfg["what"], "name": name, "frames": t["frames"],
                              "l2": "inputs larger than L2 (every frame is streamed from host memory once)"},
                   "e2e": {"value": t["frames"] / total, "unit": "frames/s",
                           "h2d_bytes_per_step": t["frames"] * frame_bytes // world, "d2h_bytes_per_step": 0},
                   "sweep": {"stream_s": stream_s, "synth_s_inside_stream": synth_s, "finish_3d_s": finish_s,
                             "merge_s": merge_s, "nccl_exchange_s": exch_s, "components": t["components"],
                             "merge": "native (csrc/api_multi.inl): ncclSend/ncclRecv of boundary planes + "
                                      "ncclAllGather of boundary-component records" if world > 1 else "single slab",
                             "spots_after_filters": t["spots_after_filters"],
                             "labelling_frames_per_s": t["frames"] / (finish_s + merge_s)},
                   "gpu_launches": None}
    else:
        r = measure_config(cfg, name, args, env, full=True)
        B, steps = r["B"], r["steps"]
        red = reduce_max([r["ms"], r["t_e2e"] * 1e3, r["k_avg"], r.get("t_e2c", 0.0) * 1e3, r["t_prod"] * 1e3])
        ms, t_e2e, k_avg, t_e2c, t_prod = red[0], red[1] / 1e3, red[2], red[3] / 1e3, red[4] / 1e3
        if rank == 0:
            alg = algorithmic_bytes_per_frame(g, dtype) * B
            achieved = alg / (k_avg * 1e-3) / 1e9
            ext = cfg["algorithm"] != "dispersion"
            ext_key = "ext_u32_dram_bytes_per_frame" if np.dtype(dtype).itemsize == 4 else "ext_u16_dram_bytes_per_frame"
            traffic = ncu_traffic_per_frame(ext_key if ext else "fused_dram_bytes_per_frame")
            h2d_gbs = r["n_e2e"] * frame_bytes / t_e2e / 1e9
            ceil_gbs, ceil_src = h2d_ceiling(world)
            out = {
                "metric": cfg["metric"], "value": world * B * steps / (ms * 1e-3), "unit": "frames/s", "n_gpus": world,
                "steps": steps, "warmup": r["warmup"], "ms_per_step": ms / steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u16" if np.dtype(dtype).itemsize == 2 else "u32",
                "data": "synthetic", "config": workload_config(cfg, name, B, r["n_distinct"]),
                "e2e": {"value": world * r["n_e2e"] / t_e2e, "unit": "frames/s", "h2d_bytes_per_step": B * frame_bytes,
                        "d2h_bytes_per_step": B * (32 + 4096 * 80), "h2d_gbs_per_gpu": h2d_gbs,
                        "h2d_ceiling_gbs_per_gpu": ceil_gbs, "h2d_ceiling_source": ceil_src,
                        "h2d_frac": (h2d_gbs / ceil_gbs) if ceil_gbs else None},
                "e2e_with_producers": {"value": world * r["n_prod_frames"] / t_prod, "unit": "frames/s",
                                       "what": "same streaming path with host threads writing every frame into the "
                                               "pinned ring inside the timed region (the decompressor's output in "
                                               "the real loop, spotfinder.cc:828-834)",
                                       "producer_threads_per_gpu": r["n_prod"]},
                "gpu_launches": int(r["launches"]),
                "roofline": {"bound": "hbm",
                             "kernel": ("ext_pass1_kernel + erode_kernel + final_kernel (summed)" if ext else
                                        "fused_dispersion_kernel (whole threshold: stream + exact resolve)"),
                             "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic * B if traffic else None, "kernel_ms": k_avg,
                             "algorithmic_bytes_per_launch": alg, "kernel_share_of_step": k_avg / (ms / steps)},
                "clocks": r.get("clocks"), "parity_check": r.get("parity_check"),
            }
            if "t_e2c" in r:
                out["e2e_from_chunks"] = {
                    "value": world * r["n_e2c"] / t_e2c, "unit": "frames/s",
                    "what": "same streaming path fed with bitshuffle-LZ4 chunks (liblz4 via pyarrow), decoded on the "
                            "device; compare with the reference arm's e2e_from_chunks",
                    "producer_threads": 4, "h2d_bytes_per_step": int(B * r["mean_chunk"]),
                    "compression_ratio": frame_bytes / r["mean_chunk"],
                    "h2d_gbs_per_gpu": r["n_e2c"] * r["mean_chunk"] / t_e2c / 1e9}
                out["e2e_compressed"] = out["e2e_from_chunks"]  # round-1 name, kept for continuity

        # ---- extra legs: the other BASELINE.json configurations, bounded -----------------------
        if name == "eiger16m_u16" and not args.no_extra:
            extra = {}
            ours_u16_us = k_avg * 1e3 / B
            frames_u16, mask_u16 = r["frames"][:4], r["mask"]
            del r
            # cfg 3: dispersion_extended, 32-bit pixels
            c3 = CONFIGS["eiger16m_u32_ext"]
            r3 = measure_config(c3, "eiger16m_u32_ext", args, env, full=False)
            red3 = reduce_max([r3["ms"], r3["t_e2e"] * 1e3, r3["k_avg"]])
            if rank == 0:
                alg3 = algorithmic_bytes_per_frame(c3["geom"], c3["dtype"]) * r3["B"]
                ach3 = alg3 / (red3[2] * 1e-3) / 1e9
                tr3 = ncu_traffic_per_frame("ext_u32_dram_bytes_per_frame")
                extra["eiger16m_u32_ext"] = {
                    "value": world * r3["B"] * r3["steps"] / (red3[0] * 1e-3), "unit": "frames/s",
                    "e2e": world * r3["n_e2e"] / (red3[1] / 1e3), "frames_per_step": r3["B"], "steps": r3["steps"],
                    "roofline": {"bound": "hbm", "kernel": "ext_pass1_kernel + erode_kernel + final_kernel (summed)",
                                 "achieved": ach3, "peak": peak, "unit": "GB/s", "frac": ach3 / peak,
                                 "kernel_ms": red3[2], "algorithmic_bytes_per_launch": alg3,
                                 "traffic": tr3 * r3["B"] if tr3 else None},
                    "parity_check": r3.get("parity_check")}
            ours_ext_us = red3[2] * 1e3 / r3["B"]
            frames_u32, mask_u32 = r3["frames"][:4], r3["mask"]
            del r3
            # cfg 5: Jungfrau 4M, stream rate + paced latency
            c5 = CONFIGS["jungfrau4m_u32"]
            r5 = measure_config(c5, "jungfrau4m_u32", args, env, full=False)
            red5 = reduce_max([r5["ms"], r5["t_e2e"] * 1e3])
            if rank == 0:
                fb5 = c5["geom"].pixels * 4
                extra["jungfrau4m_u32"] = {
                    "value": world * r5["B"] * r5["steps"] / (red5[0] * 1e-3), "unit": "frames/s",
                    "e2e": world * r5["n_e2e"] / (red5[1] / 1e3), "target_hz": 2500,
                    "h2d_gbs_per_gpu": r5["n_e2e"] * fb5 / (red5[1] / 1e3) / 1e9,
                    "parity_check": r5.get("parity_check")}
            barrier()
            if rank == 0:  # one GPU: the latency path is a single-device question
                extra["jungfrau4m_u32"]["latency_at_2500hz"] = latency_leg(c5, r5["mask"], r5["frames"], local)
            barrier()
            del r5
            # the kernel to beat
            if rank == 0:
                extra["reference_gpu"] = reference_gpu_leg({
                    "eiger16m_u16_dispersion": (frames_u16, mask_u16, "dispersion", ours_u16_us),
                    "eiger16m_u32_extended": (frames_u32, mask_u32, "dispersion_extended", ours_ext_us)})
            barrier()
            # cfg 4 at N > 1: a bounded sweep (64 frames per rank); --config sweep3600 runs the full one
            if world > 1:
                t = sweep_leg(CONFIGS["sweep3600"], env, 64 * world)
                sm = reduce_max([t["stream_s"], t["finish_s"], t["merge_s"]])
                if rank == 0:
                    extra["sweep"] = {"frames": t["frames"], "slabs": world, "stream_s": sm[0], "finish_3d_s": sm[1],
                                      "merge_s": sm[2], "components": t["components"],
                                      "note": "bounded sample (64 frames per rank); full size: --config sweep3600"}
            if rank == 0:
                out["extra"] = extra
                out["reference_gpu"] = extra.get("reference_gpu")
        if rank == 0 and world == 1 and not args.no_cpu:
            mask, frames = make_inputs(cfg, 4)
            T = cpu_threads(g)
            cpu = CpuPath(mask, T, cfg["algorithm"])
            n = T * 2
            t = cpu.run(frames, n)
            cpu.close()
            out["cpu_baseline"] = {"value": n / t, "unit": "frames/s", "cores": T, "kind": cpu.kind,
                                   "sample": f"{n} frames frame-parallel over {T} host threads; {cpu.describe()}"}
    if rank == 0:
        emit(out)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_RESULT_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its
    version banner on first use), so keep the real stdout for the result and point fd 1 at
    stderr for everything else."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, line)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=0, help="frames per step (resident in HBM); 0 = the config's own")
    ap.add_argument("--config", default="eiger16m_u16", choices=list(CONFIGS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sweep-frames", type=int, default=3600, help="--config sweep3600: frames of the whole sweep")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the other configurations' legs")
    ap.add_argument("--no-chunks", action="store_true", help="skip the compressed-ingestion legs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
