#!/bin/bash
# Sweep the bench's sort interval ("8" or "electron,ion"); one compact line per setting.
STEPS=${STEPS:-32}
for si in "$@"; do
  timeout 400 python bench.py --cells 128x128x128 --steps $STEPS --warmup 4 --sort-interval $si --no-cpu --no-e2e --no-yee 2>&1 | tail -n 1 |
    SI=$si python -c "import json,sys,os; d=json.loads(sys.stdin.read()); print('interval', os.environ['SI'], round(d['value']/1e9,2), round(d['ms_per_step'],3), round(d['roofline']['ms_per_launch'],4), round(d['sort_p']['ms_per_call'],3), d['sort_p']['calls'])"
done
