#!/bin/bash
export FFS_DEV_KNOBS=1  # the library reads its developer knobs only with this set
# A/B timing of differently-built copies of the library on ONE box (developer tool):
#   tools/ab_variants.sh name1 name2 ...   (ffs_b200/variants/libffs_b200_<name>.so; "cur" = the in-tree build)
V=fast-feedback-service_b200/ffs_b200/variants
for round in 1 2 3; do
  for n in "$@"; do
    lib=$V/libffs_b200_$n.so; [ "$n" = cur ] && lib=fast-feedback-service_b200/ffs_b200/libffs_b200.so
    r=$(FFS_B200_LIB=$PWD/$lib BATCHES=${BATCHES:-32} SLOTS=${SLOTS:-32} timeout 200 python tools/bench_threshold.py 2>&1 | tail -1)
    echo "$n: $r"
  done
done
