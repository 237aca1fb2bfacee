#!/bin/bash
# tile geometry A/B: K=4 double-buffered, K=8 double-buffered at 6 blocks/SM, K=16 at 6 blocks/SM
O=gpurun_out; mkdir -p $O
B="--steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for v in vpic_b200 vpb_k4n2 vpb_k8n2 vpb_k16 vpic_b200; do
  VPIC_B200_LIB=$PWD/vpic_b200/lib$v.so timeout 600 python bench.py $B > $O/r2aa_$v.json 2> $O/r2aa_$v.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2aa_$v.json').read().strip().splitlines()[-1])
    r=d['roofline']; print('$v', 'ms/step %.3f'%(d['ms_per_step']), r['ms_per_launch_by_species'], 'frac', r['frac'])
except Exception as e:
    print('$v failed', e); print(open('gpurun_out/r2aa_$v.err').read()[-400:])
PY
done
