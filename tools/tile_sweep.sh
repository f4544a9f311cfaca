#!/bin/bash
export FFS_DEV_KNOBS=1  # the library reads its developer knobs only with this set
# Developer sweep: rows per warp task of the fused kernel (FFS_TILE_ROWS; 0 = the library's choice).
# Waves of CTAs for a 32-frame Eiger2 16M batch = row_tiles * 0.2432 (18 strips, 4 warps per CTA, 592 CTA slots).
for tr in 0 274 222 366 182 0 274; do
 echo "tile_rows=$tr: $(FFS_TILE_ROWS=$tr BATCHES=32 SLOTS=32 python tools/bench_threshold.py 2>&1 | tail -1)"; done
