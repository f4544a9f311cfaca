#!/bin/bash
# A/B of the extended first pass on ONE box: parity tests with the variant library, then timing of
# the in-tree build and of ffs_b200/variants/libffs_b200_<name>.so   (tools/r2_ext_ab.sh ext2 ...)
export FFS_DEV_KNOBS=1
O=gpurun_out/extab; mkdir -p $O
V=$PWD/fast-feedback-service_b200/ffs_b200/variants
for n in "$@"; do
  ( FFS_B200_LIB=$V/libffs_b200_$n.so timeout 900 python -m pytest tests/test_gpu_extended.py tests/test_gpu_gain.py tests/test_gpu_trusted.py -x -q -m gpu ) > $O/pytest_$n.log 2>&1; echo "$n: $(tail -1 $O/pytest_$n.log)"
done
for round in 1 2; do
for n in cur "$@"; do
  lib=$V/libffs_b200_$n.so; [ "$n" = cur ] && lib=$PWD/fast-feedback-service_b200/ffs_b200/libffs_b200.so
  for bpp in 2 4; do
    r=$(FFS_B200_LIB=$lib BPP=$bpp ALGO=dispersion_extended BATCHES=16 SLOTS=16 timeout 300 python tools/bench_threshold.py 2>&1 | tail -1); echo "$n bpp=$bpp: $r"
  done
done
done
