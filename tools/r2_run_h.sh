#!/bin/bash
set -x
export FFS_DEV_KNOBS=1
O=gpurun_out/r2h; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x -k "extended or trusted or gain" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -4 $O/pytest.log
for w in 10 4 3; do
  echo "== ext u16 warps=$w"; FFS_EXT_WARPS=$w SLOTS=32 BATCHES=1,16,32 ALGO=dispersion_extended timeout 300 python tools/bench_threshold.py 2>&1 | tail -3
  echo "== ext u32 warps=$w"; FFS_EXT_WARPS=$w SLOTS=16 BATCHES=16 BPP=4 ALGO=dispersion_extended timeout 300 python tools/bench_threshold.py 2>&1 | tail -1
done > $O/ext_ab.log 2>&1
grep -v "^+" $O/ext_ab.log
