#!/bin/bash
# End-of-round evidence pass (run on the GPU box through gpurun).  Every step has its own timeout.
mkdir -p gpurun_out/final
O=gpurun_out/final
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.log
timeout 500 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.log
timeout 500 python bench.py --impl reference > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.log
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $O/ncu_launches.log 2>&1; echo "launches rc=$?" | tee -a $O/rc.log
BATCHES=32 SLOTS=32 timeout 400 ncu --set full --clock-control none --import-source on -k regex:fused_dispersion -s 3 -c 1 \
    -o $O/fused_full -f python tools/bench_threshold.py > $O/ncu_fused.log 2>&1; echo "fused rc=$?" | tee -a $O/rc.log
ALGO=dispersion_extended BATCHES=8 SLOTS=8 timeout 400 ncu --set full --clock-control none --import-source on -k regex:ext_pass1 -s 3 -c 1 \
    -o $O/ext_full -f python tools/bench_threshold.py > $O/ncu_ext.log 2>&1; echo "ext rc=$?" | tee -a $O/rc.log
BATCHES=32 SLOTS=32 PROCESS=1 timeout 200 python tools/bench_threshold.py > $O/thr_u16.log 2>&1
BPP=4 BATCHES=16 SLOTS=16 timeout 200 python tools/bench_threshold.py > $O/thr_u32.log 2>&1
ALGO=dispersion_extended BATCHES=8 SLOTS=8 timeout 200 python tools/bench_threshold.py > $O/thr_ext.log 2>&1
cat $O/rc.log; tail -3 $O/pytest_gpu.log; cat $O/bench.json
