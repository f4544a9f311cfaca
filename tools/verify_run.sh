#!/bin/bash
# what the driver runs at round end, on one box: GPU tests, smoke, both bench arms (timed)
export FFS_DEV_KNOBS=1
O=gpurun_out/verify; mkdir -p $O
( time timeout 2400 python -m pytest tests -x -q -m gpu ) > $O/pytest.log 2>&1; tail -4 $O/pytest.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) > $O/smoke.log 2>&1; tail -4 $O/smoke.log
( time python bench.py --impl reference ) > $O/bench_ref.json 2> $O/bench_ref.err; tail -4 $O/bench_ref.err; cut -c1-300 $O/bench_ref.json
( time python bench.py ) > $O/bench.json 2> $O/bench.err; tail -4 $O/bench.err; cut -c1-400 $O/bench.json
