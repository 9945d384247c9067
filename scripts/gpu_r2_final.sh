#!/bin/bash
# Round-2 closing run on one GPU: full GPU suite, smoke, bench (own arm + reference arm), launch list of the bench
# command and ncu captures of the kernels rewritten last.  ncu only after the same command exited 0 without it.
O=gpurun_out/${1:-r02f}; mkdir -p $O
timeout 2400 python -m pytest tests -q -m gpu > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; tail -c 600 $O/bench_n1.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference_arm.json 2> $O/bench_reference_arm.err; tail -c 400 $O/bench_reference_arm.json
B="python bench.py --sites 200000 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary"
$B > $O/plain_bench200k.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_bench_sites200k.csv $B > $O/ncu_bench200k.log 2>&1
P="python scripts/bench_blo.py 50 200000 20"
$P > $O/plain_blo_aa.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"newton_kernel|deriv_aa4" -s 20 -c 3 -o $O/newton_aa_dmma_full $P > $O/ncu_newton_aa.log 2>&1
$P > /dev/null 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_blo_aa_50x200k.csv $P > $O/ncu_blo_aa.log 2>&1
E="python bench.py --tips 65 --sites 1000000 --steps 3 --warmup 3 --no-cpu-baseline --no-secondary"
$E > $O/plain_evals.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"logl_dna4_tma" -s 4 -c 2 -o $O/logl_tma_full $E > $O/ncu_evals.log 2>&1
C="python scripts/bench_blo.py 100 20000 61"
$C > $O/plain_blo_codon.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_blo_codon_100x20k.csv $C > $O/ncu_blo_codon.log 2>&1
ls -la $O
