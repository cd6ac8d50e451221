#!/bin/bash
# A/B the LSTM kernel variants on a B200 (scratch helper).
for v in ${VARIANTS:-3 7 11 15}; do
  echo "== GNNQC_LSTM_VARIANT=$v"
  GNNQC_LSTM_VARIANT=$v python -m pytest tests/test_gpu_scoring.py -m gpu -x -q 2>&1 | tail -1
  GNNQC_LSTM_VARIANT=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value %.0f  e2e %.0f' % (d['value'], d['e2e']['value']))
for k in d['kernels']:
    print('  %-14s %.3f ms  %s' % (k['name'], k['ms_per_step'], ('%.1f TF' % k['tflops_executed']) if 'tflops_executed' in k else ''))
"
done
