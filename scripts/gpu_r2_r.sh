#!/bin/bash
# tip-table lookups with even/odd site groups fetching opposite halves (bank conflicts): A/B at C2
O=gpurun_out/r02r; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "clvs_scalers or stash_kernel or split_into or ambiguity or c2_full" > $O/pytest.log 2>&1; tail -3 $O/pytest.log
show() { python -c "
import json,sys; j=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', 'step_ms', round(j['ms_per_step'],3), 'kernel_ms', round(j['roofline']['kernel_ms'],3), j['clocks']['sm_mhz'], j['clocks']['reasons'], j['loglh'])"; }
for rep in 1 2; do
for d in 0 16; do
  PLLB200_TRAV_DEBUG=$d python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-secondary > $O/s1M_d${d}_$rep.json 2>$O/err.txt; show $O/s1M_d${d}_$rep.json
done; done
for d in 0 16; do
  PLLB200_TRAV_DEBUG=$d python bench.py --sites 200000 --steps 20 --warmup 5 --no-cpu-baseline --no-secondary > $O/s200k_d${d}.json 2>$O/err.txt; show $O/s200k_d${d}.json
done
