for cfg in "4 0" "4 75" "3 75" "4 60" "3 0" "4 0" "4 75"; do set -- $cfg
export FFS_DEV_KNOBS=1  # the library reads its developer knobs only with this set
 echo "warps/CTA=$1 carveout=$2: $(FFS_FUSED_WARPS=$1 FFS_FUSED_CARVEOUT=$2 BATCHES=32 SLOTS=32 python tools/bench_threshold.py 2>&1 | tail -1)"; done
