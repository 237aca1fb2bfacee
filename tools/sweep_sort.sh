#!/bin/bash
# Sweep the bench's sort interval; prints one compact line per setting.
for si in "$@"; do
  timeout 400 python bench.py --cells 128x128x128 --steps 16 --warmup 4 --sort-interval $si --no-cpu --no-e2e --no-yee 2>&1 | tail -n 1 |
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print('interval', $si, round(d['value']/1e9,2), round(d['ms_per_step'],3), round(d['roofline']['ms_per_launch'],4), d['sort_p'])"
done
