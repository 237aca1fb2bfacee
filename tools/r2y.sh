#!/bin/bash
# full GPU suite on the current code + cadence A/B
O=gpurun_out; mkdir -p $O
timeout 3000 python -m pytest tests -x -q -m gpu > $O/r2y_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -n 6 $O/r2y_gpu_tests.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r2y_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/r2y_smoke.log | cut -c1-200
B="--steps 24 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for si in 3,12 3,8 3,6 3,12 3,8; do
  timeout 600 python bench.py $B --sort-interval $si > $O/r2y_si$si.json 2> $O/r2y_si$si.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2y_si$si.json').read().strip().splitlines()[-1])
r=d['roofline']; print('si=$si', 'Gpush/s %.2f ms/step %.3f'%(d['value']/1e9, d['ms_per_step']), 'frac', r['frac'], r['ms_per_launch_by_species'], 'sort', round(d['sort_p']['ms_per_call'],2), d['sort_p'].get('effective'))
PY
done
