#!/bin/bash
# round-2, 8 GPUs: H2D ceilings at 1/2/4/8 concurrent GPUs, the bench line at N=8, the full cfg 4 sweep, NCCL merge test
set -x
O=gpurun_out/r2n8; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
for n in 1 2 4 8; do timeout 300 tools/ubench/h2d --gpus $n --seconds 0.7 > $O/h2d_n$n.json 2> $O/h2d_n$n.err; cat $O/h2d_n$n.json; done
python tools/make_h2d_peaks.py $O/h2d_n1.json $O/h2d_n2.json $O/h2d_n4.json $O/h2d_n8.json; cp profiles/h2d_peaks.json $O/
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -k nccl > $O/pytest_multi.log 2>&1; tail -3 $O/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_n8.json 2> $O/bench_n8.err; cat $O/bench_n8.json; tail -3 $O/bench_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --config sweep3600 > $O/sweep_n8.json 2> $O/sweep_n8.err; cat $O/sweep_n8.json; tail -3 $O/sweep_n8.err
