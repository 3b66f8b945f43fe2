#!/bin/bash
# experiment: K6 throughput on C3 against the number of CTAs (build with -DNGS_K6_GRID_ENV, library in NGS_DEALER_LIB)
for g in 148 132 111 96 74 64; do
  NGS_K6_GRID=$g timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --skip-e2e --skip-extras 2>/dev/null > /tmp/gs.json
  python -c "
import json
a=json.loads(open('/tmp/gs.json').read().strip().splitlines()[-1])
print('grid', $g, round(a['value']/1e9,2), a['parity_check']['ok'])"
done
