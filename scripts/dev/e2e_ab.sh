run() { env "$@" python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-parity --no-extras 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$*', d['roofline']['kernel_ms'], d['value'], 'e2e', d['e2e']['value'])
"; }
for rep in 1 2; do
run SAT_NOP=1
run SAT_RETIRE=0
done
