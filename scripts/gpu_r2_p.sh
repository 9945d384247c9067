#!/bin/bash
# last-round chunk balancing of the site-column kernel: an 8-GPU shard of C2 (125 k sites) on one GPU, old vs new
O=gpurun_out/r02p; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "clvs_scalers or stash_kernel or bit_stable or ambiguity or partial_traversals or spr_round or c2_full or c5_slice" > $O/pytest.log 2>&1; tail -3 $O/pytest.log
show() { python -c "
import json,sys; j=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', 'step_ms', round(j['ms_per_step'],3), 'kernel_ms', round(j['roofline']['kernel_ms'],3), 'e2e_ms', round(j['e2e']['ms_per_step'],3), j['clocks'])"; }
for s in 125000 250000 500000; do
  python bench.py --sites $s --steps 20 --warmup 5 --no-cpu-baseline --no-secondary > $O/new_$s.json 2>$O/err.txt; show $O/new_$s.json
  PLLB200_TRAV_DEBUG=8 python bench.py --sites $s --steps 20 --warmup 5 --no-cpu-baseline --no-secondary > $O/old_$s.json 2>$O/err.txt; show $O/old_$s.json
done
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-secondary > $O/new_1M.json 2>$O/err.txt; show $O/new_1M.json
PLLB200_TRAV_DEBUG=8 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-secondary > $O/old_1M.json 2>$O/err.txt; show $O/old_1M.json
