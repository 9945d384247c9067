#!/bin/bash
O=gpurun_out/r02f; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "stash_kernel or ambiguity or clvs_scalers or golden" > $O/pytest_new.log 2>&1
tail -3 $O/pytest_new.log
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/bench_$name.json 2> $O/bench_$name.err
  python - <<PY
import json
try:
    j=json.loads(open("$O/bench_$name.json").read().strip().splitlines()[-1])
    print("$name kernel_ms=%.2f ms_per_step=%.2f clocks=%s loglh=%r"%(j["roofline"]["kernel_ms"], j["ms_per_step"], j["clocks"]["sm_mhz"], j["loglh"]))
except Exception as e: print("$name failed", e)
PY
}
run base A=1
run nostore PLLB200_TRAV_DEBUG=2
run notips PLLB200_TRAV_DEBUG=1
run none PLLB200_TRAV_DEBUG=7
python bench.py --sites 200000 --steps 2 --warmup 1 --no-cpu-baseline > $O/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:trav_dna4 -s 1 -c 1 -o $O/trav_dna4_full \
    python bench.py --sites 200000 --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_trav.log 2>&1
