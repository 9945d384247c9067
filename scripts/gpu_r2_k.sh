#!/bin/bash
O=gpurun_out/r02k; mkdir -p $O
python scripts/bench_blo.py 200 1000000 > $O/blo_lean.jsonl 2>&1; tail -1 $O/blo_lean.jsonl | cut -c1-200
PLLB200_TRAV_SHORT=3 python scripts/bench_blo.py 200 1000000 > $O/blo_short3.jsonl 2>&1; tail -1 $O/blo_short3.jsonl | cut -c1-200
PLLB200_TRAV_VARIANT=3 python scripts/bench_blo.py 200 1000000 > $O/blo_old.jsonl 2>&1; tail -1 $O/blo_old.jsonl | cut -c1-200
python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-secondary > $O/bench.json 2>$O/bench.err; python -c "
import json; j=json.loads(open('$O/bench.json').read().strip().splitlines()[-1]); print('kernel_ms', j['roofline']['kernel_ms'], j['clocks'])"
