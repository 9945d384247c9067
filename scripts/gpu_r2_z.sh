#!/bin/bash
# Closing run (HEAD): full GPU suite, smoke, bench (own arm + reference arm), launch list + ncu --set full of the traversal kernel
O=gpurun_out/${1:-r02zz}; mkdir -p $O
timeout 2400 python -m pytest tests -q -m gpu > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; tail -c 200 $O/bench_n1.json; echo
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference_arm.json 2> $O/bench_reference_arm.err; tail -c 200 $O/bench_reference_arm.json; echo
export PLLB200_TRAV_SPLIT=0
B="python bench.py --sites 200000 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary"
$B > $O/plain_bench200k.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_bench_sites200k.csv $B > $O/ncu_bench200k.log 2>&1
$B > /dev/null 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trav_dna4_kernel -s 1 -c 1 -o $O/trav_dna4_full $B > $O/ncu_trav.log 2>&1
unset PLLB200_TRAV_SPLIT
S="python bench.py --sites 125000 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary"
$S > $O/plain_bench125k.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trav_dna4_split_kernel -s 2 -c 1 -o $O/trav_dna4_split_full $S > $O/ncu_trav_split.log 2>&1
ls $O
