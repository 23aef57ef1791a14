#!/bin/bash
# Developer aid: in-box A/B through environment switches (no rebuild), e.g. the activation-ring settings of the gate GEMMs
# or the specialised instantiation of the plain kernel:   gpurun -- "bash scripts/dev/env_ab.sh SAT_NO_SPEC=1"
run() { env "$@" python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-extras 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$*', d['roofline']['kernel_ms'], d['value'], 'e2e', d['e2e']['value'])
"; }
for rep in 1 2; do
run SAT_NOP=1
for v in "$@"; do run $v; done
done
