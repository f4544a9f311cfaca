#!/bin/bash
# round-2 GPU pass A: whole GPU suite with the shared-cudart build, reference-GPU timing, H2D ceiling (1 GPU)
set -x
O=gpurun_out/r2a; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -5 $O/pytest.log
timeout 300 tools/ubench/h2d --gpus 1 --seconds 1.0 > $O/h2d_n1.json 2> $O/h2d_n1.err; cat $O/h2d_n1.json
timeout 600 python tools/bench_refgpu.py > $O/refgpu.json 2> $O/refgpu.err; cat $O/refgpu.json; tail -3 $O/refgpu.err
timeout 600 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; cat $O/bench.json; tail -3 $O/bench.err
