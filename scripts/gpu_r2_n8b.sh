#!/bin/bash
# 8-GPU run: bench contract (weak + strong arm + sharding check + one smoothing pass), then the C5 SPR round sharded over the 8 GPUs
O=gpurun_out/${1:-r02w}; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 --blo > $O/bench_n8.json 2> $O/bench_n8.err; tail -c 1500 $O/bench_n8.json; tail -3 $O/bench_n8.err
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 scripts/bench_spr_c5_multi.py > $O/spr_c5_n8.json 2> $O/spr_c5_n8.err; tail -c 600 $O/spr_c5_n8.json; grep spr_round $O/spr_c5_n8.err | tail -2
