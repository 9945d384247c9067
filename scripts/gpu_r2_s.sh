#!/bin/bash
# same-box A/B of library variants: pll-modules_b200/lib/libpll_b200.<variant> copied over libpll_b200.so
O=gpurun_out/r02s; mkdir -p $O
L=pll-modules_b200/lib
show() { python -c "
import json,sys; j=json.loads(open('$1').read().strip().splitlines()[-1]); print('$1', 'step_ms', round(j['ms_per_step'],3), 'kernel_ms', round(j['roofline']['kernel_ms'],3), j['clocks']['sm_mhz'], j['clocks']['reasons'], j['loglh'])"; }
for rep in 1 2 3; do
for v in "$@"; do
  cp $L/libpll_b200.$v $L/libpll_b200.so
  python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-secondary > $O/s1M_${v}_$rep.json 2>$O/err.txt; show $O/s1M_${v}_$rep.json
done; done
for v in "$@"; do
  cp $L/libpll_b200.$v $L/libpll_b200.so
  python bench.py --sites 200000 --steps 20 --warmup 5 --no-cpu-baseline --no-secondary > $O/s200k_${v}.json 2>$O/err.txt; show $O/s200k_${v}.json
done
