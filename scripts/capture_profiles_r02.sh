#!/bin/bash
# Round-2 evidence in one GPU call: launch lists and ncu --set full captures of the kernels this round changed.
# usage (on the GPU box, from the repo root): bash scripts/capture_profiles_r02.sh
# Every ncu pass runs only after the same command exited 0 without ncu; nothing printed under ncu is a bench value.
O=gpurun_out/r02prof
mkdir -p $O
B="python bench.py --sites 200000 --steps 2 --warmup 1 --no-cpu-baseline --no-secondary"
$B > $O/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_sites200k.csv $B > $O/ncu_bench.log 2>&1
$B > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:trav_dna4 -s 1 -c 1 -o $O/trav_dna4_full $B > $O/ncu_trav.log 2>&1
# logL / derivative kernels at 1 M sites (64 taxa are enough: the kernels read two CLVs / one sumtable)
E="python bench.py --tips 64 --sites 1000000 --steps 3 --warmup 3 --no-cpu-baseline --no-secondary"
$E > $O/plain_evals.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"logl_dna|deriv_dna|sumtable_dna" -s 6 -c 12 -o $O/evals_full $E > $O/ncu_evals.log 2>&1
# protein: Newton (MODE 4), derivative pass, traversal
P="python scripts/bench_blo.py 50 200000 20"
$P > $O/plain_blo_aa.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"newton_kernel|deriv_aa4" -s 20 -c 3 -o $O/newton_aa_full $P > $O/ncu_newton_aa.log 2>&1
$P > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_blo_aa_50x200k.csv $P > $O/ncu_blo_aa.log 2>&1
# SPR round: launch list (batched candidate scoring)
S="python scripts/probes/spr_repeat.py 200 100000 1"
$S > $O/plain_spr.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_spr_200x100k.csv $S > $O/ncu_spr.log 2>&1
ls -la $O
