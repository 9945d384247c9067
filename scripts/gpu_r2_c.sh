#!/bin/bash
O=gpurun_out/r02c; mkdir -p $O
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/bench_$name.json 2> $O/bench_$name.err
  python - <<PY
import json
try:
    j=json.loads(open("$O/bench_$name.json").read().strip().splitlines()[-1])
    print("$name kernel_ms=%.2f ms_per_step=%.2f clocks=%s loglh=%r"%(j["roofline"]["kernel_ms"], j["ms_per_step"], j["clocks"]["sm_mhz"], j["loglh"]))
except Exception as e: print("$name failed", e)
PY
}
run stash0 PLLB200_TRAV_STASH=0
run tma_store PLLB200_TRAV_DEBUG=8
run tma_store_notips PLLB200_TRAV_DEBUG=9
run base_1cta PLLB200_TRAV_CTAS_PER_SM=1
run tma_1cta PLLB200_TRAV_DEBUG=8 PLLB200_TRAV_CTAS_PER_SM=1
