#!/bin/bash
O=gpurun_out/r02i; mkdir -p $O
python scripts/bench_blo.py 200 1000000 > $O/bench_blo.jsonl 2>&1; tail -2 $O/bench_blo.jsonl
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --blo ) > $O/bench_n2.json 2> $O/bench_n2.err
echo "rc=$?"; tail -c 2500 $O/bench_n2.json; tail -8 $O/bench_n2.err
