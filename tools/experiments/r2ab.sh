#!/bin/bash
# (y,z)-blocked tile order of k_advance_p: parity + A/B at full size
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > $O/r2ab_parity.log 2>&1; echo "parity rc=$?"; tail -n 5 $O/r2ab_parity.log | cut -c1-300
B="--steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for tb in 0 16,16 8,8 32,32 16,16 0 4,64; do
  VPIC_B200_TILE_BLOCK=$tb timeout 600 python bench.py $B > $O/r2ab_$tb.json 2> $O/r2ab_$tb.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2ab_$tb.json').read().strip().splitlines()[-1])
    r=d['roofline']; print('block=$tb', 'Gpush/s %.2f ms/step %.3f'%(d['value']/1e9, d['ms_per_step']), 'frac', r['frac'], r['ms_per_launch_by_species'], r['ms_by_steps_since_sort']['electron'], 'sort %.2f'%d['sort_p']['ms_per_call'])
except Exception as e:
    print('block=$tb failed', e); print(open('gpurun_out/r2ab_$tb.err').read()[-400:])
PY
done
