#!/bin/bash
# Second round-2 capture: the tile-mapped logL / derivative kernels and the protein Newton pass after their rewrite.
# Every ncu pass runs only after the same command exited 0 without ncu; nothing printed under ncu is a bench value.
O=gpurun_out/r02bprof
mkdir -p $O
E="python bench.py --tips 64 --sites 1000000 --steps 3 --warmup 3 --no-cpu-baseline --no-secondary"
$E > $O/plain_evals.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"logl_dna|deriv_dna|sumtable_dna" -s 6 -c 9 -o $O/evals_full $E > $O/ncu_evals.log 2>&1
P="python scripts/bench_blo.py 50 200000 20"
$P > $O/plain_blo_aa.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"newton_kernel|deriv_aa4" -s 20 -c 3 -o $O/newton_aa_full $P > $O/ncu_newton_aa.log 2>&1
$P > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_blo_aa_50x200k.csv $P > $O/ncu_blo_aa.log 2>&1
tail -2 $O/plain_evals.log | cut -c1-3000; cat $O/plain_blo_aa.log
ls -la $O
