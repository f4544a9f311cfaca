"""Developer tool: per-source-line executed instructions and stall samples of one kernel from an
.ncu-rep (ncu --set full --import-source on), mapped through the line table of the library the
capture ran (nvdisasm -g on its cubin).  No GPU needed.

usage: python tools/ncu_lines.py REPORT.ncu-rep LIB.so KERNEL_MANGLED_PREFIX [SOURCE_FILE_BASENAME] [min_inst]
"""
import collections, csv, io, os, re, subprocess, sys, tempfile


def main():
    rep, lib, kern = sys.argv[1:4]
    only = sys.argv[4] if len(sys.argv) > 4 else None
    min_inst = float(sys.argv[5]) if len(sys.argv) > 5 else 0.5
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr_i = next(i for i, r in enumerate(rows) if "Address" in r and "Source" in r)
    hdr, data = rows[hdr_i], rows[hdr_i + 1:]
    isamp, iex = hdr.index("# Samples"), hdr.index("Instructions Executed")
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=td, capture_output=True)
        cubin = [f for f in os.listdir(td) if f.endswith(".cubin")][0]
        dis = subprocess.run(["nvdisasm", "-g", os.path.join(td, cubin)], capture_output=True, text=True).stdout.split("\n")
    start = next(i for i, l in enumerate(dis) if l.startswith("//--------------------- .text." + kern))
    cur, line_of = None, []
    for l in dis[start + 1:]:
        if l.startswith("//---------------------"):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+\S", l):
            line_of.append(cur)
    if len(line_of) != len(data):
        print(f"warning: {len(line_of)} SASS instructions in the library, {len(data)} in the report", file=sys.stderr)
    agg = collections.defaultdict(lambda: [0, 0])
    for k, r in enumerate(data):
        if k < len(line_of) and line_of[k]:
            agg[line_of[k]][0] += int(r[isamp])
            agg[line_of[k]][1] += int(r[iex])
    tot_s = sum(v[0] for v in agg.values()) or 1
    tot_e = sum(v[1] for v in agg.values()) or 1
    srcs = {}
    print(f"total samples {tot_s}  instructions executed {tot_e}")
    for (f, ln), v in sorted(agg.items()):
        if only and f != only:
            continue
        if 100.0 * v[1] / tot_e < min_inst and 100.0 * v[0] / tot_s < min_inst:
            continue
        if f not in srcs:
            p = os.path.join(os.path.dirname(__file__), "..", "fast-feedback-service_b200", "csrc", f)
            srcs[f] = open(p).read().split("\n") if os.path.exists(p) else []
        text = srcs[f][ln - 1].strip()[:90] if ln <= len(srcs[f]) else ""
        print(f"{f}:{ln:4d}  inst {100.0*v[1]/tot_e:5.1f}%  samples {100.0*v[0]/tot_s:5.1f}%  {text}")
    other = collections.Counter()
    for (f, ln), v in agg.items():
        if only and f != only:
            other[f] += v[1]
    for f, e in other.most_common(8):
        print(f"  [{f}: inst {100.0*e/tot_e:.1f}%]")


main()
