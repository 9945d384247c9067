#!/bin/bash
# Round evidence in one GPU call: tests, bench (both arms), launch lists and ncu --set full captures.
# usage (on the GPU box, from the repo root): bash scripts/capture_profiles.sh r01
# Every ncu pass runs only after the same command exited 0 without ncu; nothing printed under ncu is a bench value.
R=${1:-r01}
O=gpurun_out/$R
mkdir -p $O
python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; tail -1 $O/pytest_gpu.log
python bench.py --steps 5 --warmup 3 --blo > $O/bench_n1.json 2> $O/bench_n1.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference_arm.json 2>> $O/bench_n1.err
python scripts/bench_configs.py c1 c3 c4 c5 > $O/bench_configs.jsonl 2> $O/bench_configs.err
python scripts/bench_blo.py 200 1000000 > $O/bench_blo.jsonl 2>&1
python scripts/bench_aa.py 500 200000 > $O/bench_aa_c3.json 2>&1
# FP64 tensor-pipe peak (the 20-state kernel's second roofline)
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/dmma_peak scripts/probes/dmma_peak.cu && /tmp/dmma_peak > $O/dmma_peak.jsonl 2>&1
# launch lists (share of the step per kernel)
python bench.py --sites 200000 --steps 2 --warmup 1 --no-cpu-baseline > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_sites200k.csv \
    python bench.py --sites 200000 --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_bench.log 2>&1
python scripts/bench_blo.py 100 1000000 > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_blo_100x1M.csv \
    python scripts/bench_blo.py 100 1000000 > $O/ncu_blo.log 2>&1
# full captures of the dominant kernels
ncu --set full --clock-control none --import-source on -k regex:trav_dna -s 1 -c 1 -o $O/trav_dna_full \
    python bench.py --sites 200000 --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_trav.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"newton|sumtable_dna" -s 40 -c 4 -o $O/blo_full \
    python scripts/bench_blo.py 50 1000000 > $O/ncu_blo_full.log 2>&1

ncu --set full --clock-control none --import-source on -k regex:trav_aa -s 1 -c 1 -o $O/trav_aa_full \
    python scripts/bench_aa.py 500 50000 > $O/ncu_trav_aa.log 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
ls -la $O
