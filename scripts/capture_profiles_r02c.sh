#!/bin/bash
# Third round-2 capture: TMA-ring kernels (protein Newton, DNA logL, derivative passes).
# Every ncu pass runs only after the same command exited 0 without ncu; nothing printed under ncu is a bench value.
O=gpurun_out/r02cprof
mkdir -p $O
P="python scripts/bench_blo.py 50 200000 20"
$P > $O/plain_blo_aa.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_blo_aa_50x200k.csv $P > $O/ncu_blo_aa.log 2>&1
$P > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"newton_aa4" -s 20 -c 3 -o $O/newton_aa_full $P > $O/ncu_newton_aa.log 2>&1
E="python bench.py --tips 65 --sites 1000000 --steps 3 --warmup 3 --no-cpu-baseline --no-secondary"
$E > $O/plain_evals.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"logl_dna4_tma|newton_dna4_tma" -s 4 -c 4 -o $O/evals_full $E > $O/ncu_evals.log 2>&1
$E > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"newton_dna4_tma" -s 2 -c 3 -o $O/deriv_full $E > $O/ncu_deriv.log 2>&1
cat $O/plain_blo_aa.log; ls -la $O
