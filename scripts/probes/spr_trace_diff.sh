#!/bin/bash
# two identical SPR rounds with every logL traced: where do they first differ?  usage: T S [option=value ...]
T=${1:-200}; S=${2:-100000}; shift; shift
PLLB200_TRACE_LOGLH=/tmp/t1.txt python scripts/probes/spr_repeat.py $T $S 1 "$@" > /dev/null
PLLB200_TRACE_LOGLH=/tmp/t2.txt python scripts/probes/spr_repeat.py $T $S 1 "$@" > /dev/null
python - "$T $S $*" <<'PY'
import sys
a=open('/tmp/t1.txt').read().split('\n'); b=open('/tmp/t2.txt').read().split('\n')
for i,(x,y) in enumerate(zip(a,b)):
    if x!=y:
        print(sys.argv[1], ": first difference at evaluation", i, "of", len(a), len(b)); break
else: print(sys.argv[1], ": identical", len(a), len(b))
PY
