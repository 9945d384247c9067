#!/bin/bash
# 8-GPU run of the bench contract (weak scaling + strong arm + sharding check)
O=gpurun_out/${1:-r02f}; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 --blo > $O/bench_n8.json 2> $O/bench_n8.err; tail -c 1200 $O/bench_n8.json; tail -3 $O/bench_n8.err
