#!/bin/bash
# advance_p shared-memory half swap A/B at full size + parity
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > $O/r2s_parity.log 2>&1; echo "parity rc=$?"; tail -n 3 $O/r2s_parity.log
B="--steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for v in vpic_b200 vpb_noswz vpic_b200 vpb_noswz; do
  VPIC_B200_LIB=$PWD/vpic_b200/lib$v.so timeout 600 python bench.py $B > $O/r2s_$v.json 2> $O/r2s_$v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2s_$v.json').read().strip().splitlines()[-1])
r=d['roofline']; print('$v', 'Gpush/s %.2f ms/step %.3f'%(d['value']/1e9, d['ms_per_step']), 'frac', r['frac'], r['ms_per_launch_by_species'], 'sort %.2f'%d['sort_p']['ms_per_call'])
PY
done
