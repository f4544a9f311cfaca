#!/bin/bash
export FFS_DEV_KNOBS=1  # the library reads its developer knobs only with this set
# Developer timing experiments on the fused dispersion kernel (results of modes != 0 are wrong by design):
#   variants/libffs_b200_dbg.so = make variant NAME=dbg EXTRA="-DFUSED_DEBUG_MODES=1"
#     FFS_FUSED_DEBUG=5 TMA traffic only; 3 streaming loop + queue pushes; 2 + stage 2 (no exact resolves); 0 everything
#   variants/libffs_b200_l1.so  = make variant NAME=l1 EXTRA="-DFUSED_EXP_L1=1": stage 3's gathers all hit a few L1 lines
V=$PWD/fast-feedback-service_b200/ffs_b200/variants
for rep in 1 2; do
  for d in 0 2 3 5; do echo "dbg=$d: $(FFS_B200_LIB=$V/libffs_b200_dbg.so FFS_FUSED_DEBUG=$d BATCHES=32 SLOTS=32 python tools/bench_threshold.py 2>&1 | tail -1)"; done
  echo "l1-hit gathers: $(FFS_B200_LIB=$V/libffs_b200_l1.so BATCHES=32 SLOTS=32 python tools/bench_threshold.py 2>&1 | tail -1)"
done
