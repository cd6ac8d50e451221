#!/bin/bash
# A/B the CML front-end (plain LDG vs TMA-staged) on a B200.
python -m pytest tests/test_gpu_scoring.py tests/test_gpu_series.py -m gpu -x -q 2>&1 | tail -2
for fe in plain tma; do
  echo "== GNNQC_FRONTEND=$fe"
  GNNQC_FRONTEND=$fe python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value %.0f  e2e %.0f' % (d['value'], d['e2e']['value']))
for k in d['kernels']:
    if 'frontend' in k['name'] or 'coef' in k['name']: print('  %-14s %.3f ms' % (k['name'], k['ms_per_step']))
"
done
