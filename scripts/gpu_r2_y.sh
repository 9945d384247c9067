#!/bin/bash
# usage: gpu_r2_y.sh <variant under test> — parity subset with that variant, then same-box A/B against libpll_b200.base
O=gpurun_out/r02y; mkdir -p $O
V=${1:-tsm}
cp pll-modules_b200/lib/libpll_b200.$V pll-modules_b200/lib/libpll_b200.so
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or clvs_scalers or stash_kernel or split_into or bit_stable or ambiguity or spr_round or partial_trav or c2_full or c5_slice or multi_partition" > $O/pytest.log 2>&1; tail -3 $O/pytest.log
bash scripts/gpu_r2_s.sh base $V
