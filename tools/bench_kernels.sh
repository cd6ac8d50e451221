#!/bin/bash
# Scoring parity tests, then the default bench with the per-kernel table (CUDA-event hooks).
timeout 300 python -m pytest tests/test_gpu_scoring.py -m gpu -x -q 2>&1 | tail -4
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" 2>gpurun_out/bench.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value %.0f  e2e %.0f' % (d['value'], d['e2e']['value']))
for k in d['kernels']: print('  %-14s %.3f ms' % (k['name'], k['ms_per_step']))"
tail -3 gpurun_out/bench.err
