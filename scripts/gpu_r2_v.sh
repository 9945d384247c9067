#!/bin/bash
# sub-list traversal with the top fused into the same launch (tickets): parity + timing at shard sizes of C2
O=gpurun_out/r02v; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "split_into or stash_kernel or bit_stable or c5_slice or spr_round" > $O/pytest.log 2>&1; tail -3 $O/pytest.log
show() { python -c "
import json,sys; j=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', 'step_ms', round(j['ms_per_step'],3), 'kernel_ms', round(j['roofline']['kernel_ms'],3), 'e2e_ms', round(j['e2e']['ms_per_step'],3), 'launches', j['gpu_launches'], j['clocks']['sm_mhz'], j['clocks']['reasons'], j['loglh'])"; }
export PLLB200_TRAV_SPLIT_ROUNDS=100000
for s in 125000 250000; do
  for t in 0 64 128 256; do
    PLLB200_TRAV_SPLIT=$t python bench.py --sites $s --steps 20 --warmup 5 --no-cpu-baseline --no-secondary > $O/s${s}_t$t.json 2>$O/err.txt; show $O/s${s}_t$t.json
  done
done
for t in 0 128 256; do
  PLLB200_TRAV_SPLIT=$t python bench.py --sites 500000 --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > $O/s500000_t$t.json 2>$O/err.txt; show $O/s500000_t$t.json
done
for t in 0 128 256; do
  PLLB200_TRAV_SPLIT=$t python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-secondary > $O/s1M_t$t.json 2>$O/err.txt; show $O/s1M_t$t.json
done
