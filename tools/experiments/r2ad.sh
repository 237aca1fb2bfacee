#!/bin/bash
# lazy page-locking in RESIDENT mode + no copy of stale host particles on resize: drop-in tests, deck traces, deck_e2e
O=$PWD/gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_dropin.py -x -q > $O/r2ad_dropin.log 2>&1; echo "dropin rc=$?"; tail -n 4 $O/r2ad_dropin.log | cut -c1-300
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q > $O/r2ad_parity.log 2>&1; echo "parity rc=$?"; tail -n 2 $O/r2ad_parity.log | cut -c1-300
bash tools/experiments/r2ac.sh 2>/dev/null
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 > $O/r2ad_bench.json 2> $O/r2ad_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r2ad_bench.json').read().strip().splitlines()[-1])
for k in ('harris_c3','lpi_c5'):
    x=d['deck_e2e'][k]; print(k, {a: x[a] for a in x if a.startswith('speedup')}, 'gpu wall', x['gpu']['wall_s'], x['gpu']['stages_s'], 'cpu wall', x['cpu']['wall_s'])
PY
