#!/bin/bash
# Developer aid: A/B two builds of libsat.so on ONE GPU box (boxes of the pool differ by ~2 % and clocks drift between
# processes, so two gpurun calls cannot resolve a 1 % change).  Put the builds at build_ab/libsat_old.so and
# build_ab/libsat_new.so (git-ignored, shipped to the box), then:  gpurun -- 'bash scripts/dev/ab_bench.sh'
for rep in 1 2; do for v in old new; do
SAT_LIB=$PWD/build_ab/libsat_$v.so python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-extras 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$v', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value'], d['clocks']['sm_mhz'])
"; done; done
