#!/bin/bash
# final pass of round 2 (1 GPU): bench lines, launch list, --set full captures of the kernels that changed
# after tools/r2_profile.sh ran (erosion, final pass, decoder).  gpurun merges at most 64 MiB back:
# every report is summarised on the box and only the JSON is kept.
set -x
export FFS_DEV_KNOBS=1
O=gpurun_out/r2c; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on"
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; cut -c1-300 $O/bench_n1.json; tail -3 $O/bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cut -c1-200 $O/bench_ref.json
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > $O/b2.json 2> $O/b2.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches_bench_steps2.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > $O/ncu_launches.log 2>&1
ALGO=dispersion_extended BATCHES=16 SLOTS=16 timeout 600 $NCU -k regex:erode_kernel -s 3 -c 1 -o $O/erode python tools/bench_threshold.py > $O/n4.log 2>&1
ALGO=dispersion_extended BATCHES=16 SLOTS=16 timeout 600 $NCU -k regex:final_kernel -s 3 -c 1 -o $O/final python tools/bench_threshold.py > $O/n5.log 2>&1
timeout 600 $NCU -k regex:bslz4_decode -s 2 -c 1 -o $O/bslz4 python tools/bench_bslz4.py > $O/n6.log 2>&1
for k in erode final bslz4; do python tools/summarize_ncu.py $O/$k.ncu-rep $O/${k}_ncu_full.json "tools/r2c_profile.sh" && rm -f $O/$k.ncu-rep; done
ls -la $O
