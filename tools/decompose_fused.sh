L=$PWD/fast-feedback-service_b200/ffs_b200/variants/libffs_b200_dbg.so
for d in 0 2 3 5 0 2 3 5; do echo "dbg=$d: $(FFS_B200_LIB=$L FFS_FUSED_DEBUG=$d BATCHES=32 SLOTS=32 python tools/bench_threshold.py 2>&1 | tail -1)"; done
for bg in 0.5 10 50; do echo "bg=$bg dbg=0: $(FFS_B200_LIB=$L BG=$bg BATCHES=32 SLOTS=32 python tools/bench_threshold.py 2>&1 | tail -1)"; echo "bg=$bg dbg=3: $(FFS_B200_LIB=$L BG=$bg FFS_FUSED_DEBUG=3 BATCHES=32 SLOTS=32 python tools/bench_threshold.py 2>&1 | tail -1)"; done
