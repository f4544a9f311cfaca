#!/usr/bin/env python
"""bench.py -- frames/s of the DIALS dispersion spot-finding hot path on B200.

Workload (BASELINE.json configs[1]): Eiger2 16M (4148x4362) 16-bit frames, standard
dispersion (7x7, nsig_b 6, nsig_s 3, min_count 2) + 2D connected components + per-spot
reduction, synthetic Poisson background + Gaussian spots + module-gap mask, frames sharded
frame-wise across ranks (no data-path collective: "scaling": "weak").

  value : whole-job frames/s with the frames already resident in HBM (slot API:
          ffs_process on a batch of B frames per step; B*36 MB is far larger than L2).
  e2e   : the same metric through the streaming C-ABI calls (ffs_acquire / ffs_submit /
          ffs_wait) with HOST buffers: every frame is copied from pinned host memory to the
          device inside the timed region and its counters + spot list are read back.
  roofline    : the threshold prefilter kernel, bracketed by CUDA events on its own stream
                inside the timed region; algorithmic bytes per frame = image + 1-bit validity
                plane + 1-bit strong plane.
  cpu_baseline: the reference's CPU path on this box's host cores (rank 0, N=1, bounded
                sample): threshold = the reference's standalone.cc compiled in place
                (oracle/_ref), connected components = the oracle port.

`--impl reference` times that CPU path alone with all host threads (rank 0 only).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "fast-feedback-service_b200"))

from ffs_b200 import synth  # noqa: E402

METRIC = "frames/sec (Eiger2 16M 16-bit)"
GEOM = synth.EIGER2_16M
DTYPE = np.uint16
N_DISTINCT = 8  # distinct synthetic frames (generation costs ~1 s each); cycled


def algorithmic_bytes_per_frame(g=GEOM, dtype=DTYPE) -> int:
    return g.pixels * np.dtype(dtype).itemsize + 2 * ((g.pixels + 7) // 8)


def make_inputs(n_distinct=N_DISTINCT):
    mask = synth.make_mask(GEOM)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        frames = list(ex.map(lambda i: synth.make_frame(GEOM, mask, DTYPE, 2, i), range(n_distinct)))
    return mask, np.stack(frames)


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU implementation of the path
# ---------------------------------------------------------------------------------------------
class CpuPath:
    """threshold: reference standalone.cc (oracle/_ref) when built, else the oracle port;
    connected components + filters: the oracle port (Boost.Graph CC is not buildable here)."""

    def __init__(self, mask, n_threads):
        from oracle import oracle
        self.o = oracle
        self.mask = np.ascontiguousarray(mask, np.uint8)
        self.kind = "reference" if oracle.ref_available() else "port"
        self.n_threads = n_threads
        self.h, self.w = mask.shape
        self.local = threading.local()
        self.L = oracle.lib()
        self.R = oracle.ref() if self.kind == "reference" else None
        # ONE pool for the object's lifetime: a reference context is ~0.6 GB at 16M and lives in
        # its worker thread, so a pool per run() would leak n_threads contexts per step.
        self.ex = ThreadPoolExecutor(max_workers=n_threads)
        self.ctxs = []
        self.lock = threading.Lock()

    def _ctx(self):
        if not hasattr(self.local, "ctx"):
            self.local.dst = np.empty((self.h, self.w), np.uint8)
            self.local.ctx = self.R.ffs_ref_create(self.w, self.h) if self.R else None
            if self.local.ctx:
                with self.lock:
                    self.ctxs.append(self.local.ctx)
        return self.local.ctx, self.local.dst

    def close(self):
        self.ex.shutdown(wait=True)
        for c in self.ctxs:
            self.R.ffs_ref_free(c)
        self.ctxs = []

    def one(self, frame):
        ctx, dst = self._ctx()
        if self.R:
            self.R.ffs_ref_standard_dispersion(ctx, frame.ctypes.data, frame.dtype.itemsize,
                                               self.mask.ctypes.data, dst.ctypes.data)
        else:
            p = self.o.default_params()
            self.L.ffs_oracle_dispersion(frame.ctypes.data, frame.dtype.itemsize, self.mask.ctypes.data,
                                         self.w, self.h, C.byref(p), dst.ctypes.data)
        _, spots = self.o.connected_components(dst, frame, row_wrap=True, want_labels=False)
        kept, _, _ = self.o.filter_spots(spots, 3, 2.0)
        return len(kept)

    def run(self, frames, n_frames):
        """Process n_frames (cycled) frame-parallel over n_threads; returns seconds."""
        ex = self.ex
        if len(self.ctxs) < self.n_threads:  # first call: let every worker build its context
            list(ex.map(self.one, [frames[i % len(frames)] for i in range(self.n_threads)]))
        t0 = time.perf_counter()
        list(ex.map(self.one, [frames[i % len(frames)] for i in range(n_frames)]))
        return time.perf_counter() - t0


def cpu_threads() -> int:
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    try:  # ~0.6 GB per reference context at 16M (24-byte SAT cell + double image)
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
        n = min(n, max(1, int(avail * 0.5 / 0.7e9)))
    except Exception:
        pass
    return max(1, min(n, 64))


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    mask, frames = make_inputs(4)
    T = cpu_threads()
    cpu = CpuPath(mask, T)
    per_step = T  # one frame per thread per step: a bounded sample of the stream
    for _ in range(args.warmup):
        cpu.run(frames, per_step)
    t = sum(cpu.run(frames, per_step) for _ in range(args.steps))
    cpu.close()
    fps = per_step * args.steps / t
    sample = (f"{per_step} frames/step x {args.steps} steps, frame-parallel over {T} threads; threshold = "
              f"{'reference standalone.cc compiled in place' if cpu.kind == 'reference' else 'oracle port'}, "
              "CC + filters = oracle port")
    emit(({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(per_step),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": T, "kind": cpu.kind, "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(batch):
    return {"workload": "Eiger2 16M (4148x4362) u16: dispersion 7x7 nsig_b=6 nsig_s=3 min_count=2 + 2D connected "
                        "components + spot reduction; synthetic Poisson(2)+Gaussian spots, module-gap mask; "
                        "frame-wise sharding",
            "frames_per_step": batch, "l2": "inputs larger than L2 (batch of frames per step >> 126 MB)",
            "distinct_frames": N_DISTINCT}


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


def ncu_traffic_per_frame():
    """dram bytes (read+write) per frame of the prefilter kernel from the committed ncu capture."""
    p = ROOT / "profiles" / "roofline_traffic.json"
    try:
        return float(json.loads(p.read_text())["fused_dram_bytes_per_frame"])
    except Exception:
        return None


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    from ffs_b200.spotfinder import Spotfinder

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch
    mask, frames = make_inputs()
    sf = Spotfinder(GEOM.width, GEOM.height, DTYPE, mask, n_slots=B, device=local)
    for s0 in range(0, B, len(frames)):
        sf.upload(frames[: min(len(frames), B - s0)], s0)
    stream = torch.cuda.Stream()
    sh = stream.cuda_stream

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: HBM-resident frames through the slot API --------------------------------------
    for _ in range(max(args.warmup, 3)):
        sf.process(0, B, sh, sync=True)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    sf.profile_enable(True)
    l0 = sf.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        sf.process(0, B, sh, sync=False)
    e1.record(stream)
    e1.synchronize()
    barrier()
    launches = sf.kernel_launches - l0
    ms = e0.elapsed_time(e1)
    k_ms, k_n = sf.profile_read()
    sf.profile_enable(False)
    clocks = sampler.stop() if rank == 0 else None
    res, _ = sf.fetch(0, 16)
    assert res.status == 0 and res.num_strong_pixels > 0

    # ---- e2e: streaming API, pinned host frames -> H2D -> kernels -> results on host ----------
    ring = min(B, 8)
    sf.close()
    sf = Spotfinder(GEOM.width, GEOM.height, DTYPE, mask, n_slots=ring, device=local)
    n_e2e = B * args.steps

    def stream_frames(n, fill=False):
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

    stream_frames(ring * 2, fill=True)  # warm-up, and fills the pinned ring (outside the timed region)
    barrier()
    t0 = time.perf_counter()
    last = stream_frames(n_e2e)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    assert last.status == 0 and last.num_strong_pixels > 0
    barrier()

    # ---- e2e, compressed ingestion: the raw bitshuffle-LZ4 chunk (what the detector writes and
    #      the reference's reader returns) crosses PCIe and is decoded on the device -------------
    from ffs_b200 import bslz4_encode
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        chunks = list(ex.map(bslz4_encode.compress_chunk, frames))
    chunk_len = {}

    def stream_chunks(n, fill=False):
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

    stream_chunks(ring * 2, fill=True)  # warm-up; every ring slot now holds a chunk
    # Producer threads as in the reference's frame loop (spotfinder.cc:725: N reader threads share
    # the ring): each indexes its chunk's block headers and submits; this thread collects results.
    n_prod = 4
    n_e2c = (n_e2e * 4 // n_prod) * n_prod

    def producer(k):
        for i in range(n_e2c // n_prod):
            slot, _ = sf.acquire_compressed()
            sf.submit_compressed(slot, k + i * n_prod, chunk_len[slot])

    barrier()
    threads = [threading.Thread(target=producer, args=(k,)) for k in range(n_prod)]
    t0 = time.perf_counter()
    for t in threads:
        t.start()
    got, last_c = 0, None
    while got < n_e2c:
        if sf.pending:
            last_c, _ = sf.wait(16)
            got += 1
        else:
            time.sleep(2e-5)
    torch.cuda.synchronize()
    t_e2c = time.perf_counter() - t0
    for t in threads:
        t.join()
    assert last_c.status == 0 and last_c.num_strong_pixels > 0
    mean_chunk = float(np.mean([len(c) for c in chunks]))
    barrier()

    # ---- reduce over ranks: max time -------------------------------------------------------
    if world > 1:
        t = torch.tensor([ms, t_e2e * 1e3, k_ms / max(k_n, 1), t_e2c * 1e3], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, t_e2e, k_avg, t_e2c = float(t[0]), float(t[1]) / 1e3, float(t[2]), float(t[3]) / 1e3
    else:
        k_avg = k_ms / max(k_n, 1)
    if rank == 0:
        frame_bytes = GEOM.pixels * np.dtype(DTYPE).itemsize
        peak, peak_src = measured_peak_hbm()
        alg = algorithmic_bytes_per_frame() * B
        achieved = alg / (k_avg * 1e-3) / 1e9
        traffic = ncu_traffic_per_frame()
        out = {
            "metric": METRIC, "value": world * B * args.steps / (ms * 1e-3), "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u16",
            "data": "synthetic", "config": workload_config(B),
            "e2e": {"value": world * n_e2e / t_e2e, "unit": "frames/s", "h2d_bytes_per_step": B * frame_bytes,
                    "d2h_bytes_per_step": B * (32 + 4096 * 80),
                    "h2d_gbs_per_gpu": n_e2e * frame_bytes / t_e2e / 1e9},
            "e2e_compressed": {"value": world * n_e2c / t_e2c, "unit": "frames/s",
                               "what": "same streaming path fed with bitshuffle-LZ4 chunks (liblz4 via pyarrow), "
                                       "decoded on the device; NOT comparable with the reference arm, which starts "
                                       "from dense frames",
                               "producer_threads": n_prod,
                               "h2d_bytes_per_step": int(B * mean_chunk), "compression_ratio": frame_bytes / mean_chunk,
                               "h2d_gbs_per_gpu": n_e2c * mean_chunk / t_e2c / 1e9},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "fused_dispersion_kernel<u16> (whole threshold: stream + exact resolve)", "achieved": achieved, "peak": peak,
                         "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic * B if traffic else None, "kernel_ms": k_avg,
                         "algorithmic_bytes_per_launch": alg,
                         "kernel_share_of_step": k_avg / (ms / args.steps)},
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            T = cpu_threads()
            cpu = CpuPath(mask, T)
            n = T * 2
            t = cpu.run(frames, n)
            cpu.close()
            out["cpu_baseline"] = {"value": n / t, "unit": "frames/s", "cores": T, "kind": cpu.kind,
                                   "sample": f"{n} frames frame-parallel over {T} host threads; threshold = "
                                             f"{'reference standalone.cc (oracle/_ref)' if cpu.kind == 'reference' else 'oracle port'}"
                                             ", CC + filters = oracle port"}
        emit(out)
    sf.close()
    if world > 1:
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
    ap.add_argument("--batch", type=int, default=32, help="frames per step (resident in HBM)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
