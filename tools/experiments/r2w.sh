#!/bin/bash
# where the electron launch's time goes: builds that skip the crossers' / strays' / runs' reductions (timing only)
O=gpurun_out; mkdir -p $O
B="--steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for v in vpic_b200 vpb_x1 vpb_x2 vpb_x4 vpb_x7 vpic_b200; do
  VPIC_B200_LIB=$PWD/vpic_b200/lib$v.so timeout 600 python bench.py $B > $O/r2w_$v.json 2> $O/r2w_$v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2w_$v.json').read().strip().splitlines()[-1])
r=d['roofline']; print('$v', 'ms/step %.3f'%(d['ms_per_step']), r['ms_per_launch_by_species'], r['ms_by_steps_since_sort'])
PY
done
