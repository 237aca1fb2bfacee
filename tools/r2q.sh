#!/bin/bash
# sort_p ILP rewrite: parity + 1-GPU bench (reduced legs) + sort launch list at 128^3
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > $O/r2q_parity.log 2>&1; echo "parity rc=$?"; tail -n 3 $O/r2q_parity.log
timeout 900 python -m pytest tests/test_c4_full_size_properties.py -x -q -m gpu > $O/r2q_c4.log 2>&1; echo "c4 rc=$?"; tail -n 3 $O/r2q_c4.log
timeout 900 python bench.py --steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks --breakdown 6 > $O/r2q_bench.json 2> $O/r2q_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r2q_bench.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches')}); print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch_by_species'], 'sort', d['sort_p'])
print('breakdown', d.get('breakdown_ms'))
PY
C128="python bench.py --cells 128x128x128 --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_sort|k_hist|k_scan' -c 60 --csv --log-file $O/r2q_sort_launches_128.csv $C128 > $O/r2q_ncu.log 2>&1; echo "ncu rc=$?"
grep -c k_sort $O/r2q_sort_launches_128.csv
