#!/bin/bash
# full GPU suite (all failures reported), then the bench line with its secondary C3/C4/C5 keys
O=gpurun_out/${1:-r02s}_tests; mkdir -p $O
timeout 2400 python -m pytest tests -q -m gpu > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log; tail -15 $O/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; tail -c 3000 $O/bench_n1.json
