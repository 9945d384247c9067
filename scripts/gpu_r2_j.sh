#!/bin/bash
O=gpurun_out/r02j; mkdir -p $O
python scripts/bench_blo.py 100 1000000 > $O/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_blo_100x1M.csv \
    python scripts/bench_blo.py 100 1000000 > $O/ncu_blo.log 2>&1
PLLB200_TRAV_VARIANT=3 python scripts/bench_blo.py 200 1000000 > $O/bench_blo_old.jsonl 2>&1; tail -1 $O/bench_blo_old.jsonl
python scripts/launch_summary.py $O/launches_blo_100x1M.csv | head -20
