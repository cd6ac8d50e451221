#!/bin/bash
# One short bench line per workload (device-resident value and e2e), for the README table.
for wl in cml_gcn cml_baseline soilnet_gcn cml_series ig train; do
  extra=""
  [ "$wl" = "ig" ] && extra="--steps 2 --warmup 1"
  [ "$wl" != "ig" ] && extra="--steps 5 --warmup 3"
  timeout 300 python bench.py --workload $wl $extra --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('%-14s value %12.0f %s   e2e %12.0f   ms/step %.3f  launches %s' % ('$wl', d['value'], d['unit'], d['e2e']['value'], d['ms_per_step'], d.get('gpu_launches')))
for k in d.get('kernels', [])[:4]: print('      %-18s %.3f ms' % (k['name'], k['ms_per_step']))"
done
