#!/bin/bash
export FFS_DEV_KNOBS=1  # the library reads its developer knobs only with this set
# Developer sweep: TMA pipeline depth x warps per CTA of the fused kernel (FFS_FUSED_STAGES / FFS_FUSED_WARPS).
for cfg in "2 4" "3 4" "3 3" "3 2" "2 4" "3 4"; do set -- $cfg
 echo "stages=$1 warps/CTA=$2: $(FFS_FUSED_STAGES=$1 FFS_FUSED_WARPS=$2 BATCHES=32 SLOTS=32 python tools/bench_threshold.py 2>&1 | tail -1)"; done
