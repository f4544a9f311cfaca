#!/bin/bash
# A/B of the device bitshuffle-LZ4 decoder on ONE box: parity tests on the in-tree build, then
# kernel time (ncu gpu__time_duration) of the in-tree build and of ffs_b200/variants/libffs_b200_<name>.so
#   tools/r2_bslz4_ab.sh base lz1 ...
export FFS_DEV_KNOBS=1
O=gpurun_out/bslz4ab; mkdir -p $O
V=$PWD/fast-feedback-service_b200/ffs_b200/variants
( time timeout 900 python -m pytest tests/test_gpu_bslz4.py tests/test_gpu_cli.py -x -q -m gpu ) > $O/pytest.log 2>&1; tail -5 $O/pytest.log
M=gpu__time_duration.sum,smsp__inst_executed.sum,sm__cycles_active.avg,sm__cycles_elapsed.max,smsp__issue_active.avg.per_cycle_active
for n in cur "$@"; do
  lib=$V/libffs_b200_$n.so; [ "$n" = cur ] && lib=$PWD/fast-feedback-service_b200/ffs_b200/libffs_b200.so
  for bpp in 2 4; do
    r=$(FFS_B200_LIB=$lib BPP=$bpp timeout 300 python tools/bench_bslz4.py 2>&1 | tail -1); echo "$n bpp=$bpp: $r"
    FFS_B200_LIB=$lib BPP=$bpp timeout 300 ncu --metrics $M --clock-control none -k regex:bslz4_decode -s 2 -c 3 --csv --log-file $O/ncu_${n}_$bpp.csv python tools/bench_bslz4.py > $O/ncu_${n}_$bpp.log 2>&1
    grep -E "bslz4_decode" $O/ncu_${n}_$bpp.csv | awk -F'","' '{print $(NF-2), $NF}' | tr -d '"' | paste -sd' ' | cut -c1-400
  done
done
