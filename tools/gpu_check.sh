#!/bin/bash
# run on the GPU box: parity tests, then a short bench; prints one summary line
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 3 --warmup 3 --no-cpu --no-queries "$@" 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('GB/s %.1f  ms %.2f  e2e %s  stages %s' % (d['value'], d['ms_per_step'], d['e2e'] and round(d['e2e']['value'],1), d['roofline']['stages_ms']))"
