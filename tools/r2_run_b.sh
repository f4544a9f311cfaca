#!/bin/bash
# round-2 GPU pass B: full GPU suite, new bench line (all legs), refgpu timing
set -x
export FFS_DEV_KNOBS=1
O=gpurun_out/r2b; mkdir -p $O
timeout 1700 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -15 $O/pytest.log
timeout 600 python tools/bench_refgpu.py > $O/refgpu.json 2> $O/refgpu.err; cat $O/refgpu.json; tail -3 $O/refgpu.err
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; cat $O/bench.json; tail -5 $O/bench.err
FFS_H2D_2D=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu --no-chunks > $O/bench_2d.json 2> $O/bench_2d.err; cat $O/bench_2d.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cat $O/bench_ref.json; tail -3 $O/bench_ref.err
