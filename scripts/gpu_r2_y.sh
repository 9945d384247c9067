#!/bin/bash
O=gpurun_out/r02y; mkdir -p $O
cp pll-modules_b200/lib/libpll_b200.ta2 pll-modules_b200/lib/libpll_b200.so
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or clvs_scalers or stash_kernel or split_into or bit_stable or ambiguity or spr_round or partial_trav" > $O/pytest.log 2>&1; tail -3 $O/pytest.log
bash scripts/gpu_r2_s.sh ta1 ta2
