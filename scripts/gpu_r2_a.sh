#!/bin/bash
# round 2, call A: lean DNA kernel — parity first, then bench old vs new
mkdir -p gpurun_out/r02a
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "stash_kernel or ambiguity or clvs_scalers or golden" > gpurun_out/r02a/pytest_new.log 2>&1
echo "pytest_new rc=$?" >> gpurun_out/r02a/pytest_new.log
tail -5 gpurun_out/r02a/pytest_new.log
PLLB200_TRAV_VARIANT=3 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02a/bench_old.json 2> gpurun_out/r02a/bench_old.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02a/bench_new.json 2> gpurun_out/r02a/bench_new.err
PLLB200_TRAV_STASH=0 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02a/bench_new_stash0.json 2> gpurun_out/r02a/bench_new_stash0.err
tail -c 600 gpurun_out/r02a/bench_old.json; echo; tail -c 1500 gpurun_out/r02a/bench_new.json; echo; tail -c 600 gpurun_out/r02a/bench_new_stash0.json
tail -3 gpurun_out/r02a/bench_new.err
