#!/bin/bash
O=gpurun_out/r02w; mkdir -p $O
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29524 bench.py --gpus 4 --steps 5 --warmup 3 > $O/bench_n4.json 2> $O/bench_n4.err; tail -c 300 $O/bench_n4.json; tail -2 $O/bench_n4.err
