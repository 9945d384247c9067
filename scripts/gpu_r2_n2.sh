#!/bin/bash
# 2-GPU run of the bench contract (weak + strong scaling, sharding check) and the multi-rank GPU tests
O=gpurun_out/${1:-r02f}; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; tail -c 1500 $O/bench_n2.json; tail -3 $O/bench_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --blo --no-strong > $O/bench_n2_blo.json 2> $O/bench_n2_blo.err; tail -c 600 $O/bench_n2_blo.json
