#!/bin/bash
O=gpurun_out/r02h; mkdir -p $O
( time timeout 900 python bench.py --steps 5 --warmup 3 ) > $O/bench_n1.json 2> $O/bench_n1.err
tail -c 3000 $O/bench_n1.json; tail -5 $O/bench_n1.err
