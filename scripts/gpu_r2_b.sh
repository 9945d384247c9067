#!/bin/bash
# round 2, call B: ablations of the lean DNA kernel (timing only) + ncu full capture at 200k sites
O=gpurun_out/r02b; mkdir -p $O
for d in 0 1 2 4 3 7; do
  PLLB200_TRAV_DEBUG=$d timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/bench_dbg$d.json 2> $O/bench_dbg$d.err
  python - <<PY
import json
try:
    j=json.loads(open("$O/bench_dbg$d.json").read().strip().splitlines()[-1])
    print("dbg=$d kernel_ms=%.2f ms_per_step=%.2f clocks=%s"%(j["roofline"]["kernel_ms"], j["ms_per_step"], j["clocks"]["sm_mhz"]))
except Exception as e: print("dbg=$d failed", e)
PY
done
python bench.py --sites 200000 --steps 2 --warmup 1 --no-cpu-baseline > $O/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:trav_dna4 -s 1 -c 1 -o $O/trav_dna4_full \
    python bench.py --sites 200000 --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_trav.log 2>&1
ls -la $O
