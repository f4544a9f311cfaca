"""Turn an .ncu-rep (gpurun_out/, scratch) into the small JSON summary committed under profiles/.

usage: python tools/summarize_ncu.py gpurun_out/X.ncu-rep profiles/rN_X.json "<command that produced it>"

Reads the report with `ncu -i ... --page raw --csv` (per-launch metrics) and
`--page source --csv` (per-instruction stall samples; needs -lineinfo) -- no GPU needed.
"""
import csv
import io
import json
import subprocess
import sys

KEEP = [
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__issue_active.avg.per_cycle_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tma.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
]


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, dst, cmd = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
    raw = ncu_csv(rep, "raw")
    hdr, units = raw[0], raw[1]
    launches = []
    for r in raw[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")]}
        for h, u, v in zip(hdr, units, r):
            if h in KEEP or (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")):
                d[h] = [v, u]
        launches.append(d)
    summary = {"report": rep, "command": cmd, "launches": launches}
    src = ncu_csv(rep, "source")
    # first kernel's table only: header row is the one starting with "Address"
    try:
        h_i = next(i for i, r in enumerate(src) if r and r[0] == "Address")
        shdr = src[h_i]
        body = []
        for r in src[h_i + 1:]:
            if not r or r[0] in ("Kernel Name", "Address"):
                break
            body.append(r)
        i_src, i_s, i_e = shdr.index("Source"), shdr.index("# Samples"), shdr.index("Instructions Executed")
        stall = [(i, h[6:]) for i, h in enumerate(shdr) if h.startswith("stall_") and "Not Issued" not in h]
        tot = sum(int(r[i_s] or 0) for r in body) or 1
        bins = []
        for b in range(0, len(body), 40):
            seg = body[b:b + 40]
            s = sum(int(r[i_s] or 0) for r in seg)
            if s < tot * 0.005:
                continue
            st = sorted(((sum(int(r[c] or 0) for r in seg), n) for c, n in stall), reverse=True)[:3]
            bins.append({"sass_index": b, "samples_pct": round(100.0 * s / tot, 2),
                         "inst_executed": sum(int(r[i_e] or 0) for r in seg),
                         "top_stalls": [[n, v] for v, n in st], "first_inst": seg[0][i_src].strip()[:60]})
        summary["source_hotspots_40_instruction_bins"] = {"total_samples": tot, "sass_instructions": len(body), "bins": bins}
    except StopIteration:
        pass
    with open(dst, "w") as f:
        json.dump(summary, f, indent=1)
    print("wrote", dst, "launches", len(launches))


if __name__ == "__main__":
    main()
