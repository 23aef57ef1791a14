#!/bin/bash
# Developer aid: in-box A/B of the activation-ring settings of the gate GEMMs through environment switches (no rebuild):
#   gpurun -- "bash scripts/dev/env_ab.sh"
run() { env "$@" python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-extras 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$*', d['roofline']['kernel_ms'], d['value'])
"; }
for rep in 1 2; do
run SAT_NOP=1
run SAT_TC_G=2 SAT_TC_STAGES=4
run SAT_TC_G=2 SAT_TC_STAGES=3
run SAT_TC_PREFILL=1
run SAT_TC_PREFILL=0
done
