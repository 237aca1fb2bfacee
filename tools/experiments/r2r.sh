#!/bin/bash
# tracked-test transfer trace + advance_p record-prefetch A/B at full size
O=gpurun_out; mkdir -p $O
VPIC_B200_TRACE_TRANSFERS=1 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -s -k tracked > $O/r2r_tracked_alone.log 2>&1; echo "tracked alone rc=$?"
VPIC_B200_TRACE_TRANSFERS=1 timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > $O/r2r_parity_all.log 2>&1; echo "parity all rc=$?"; tail -n 3 $O/r2r_parity_all.log
B="--steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for v in vpic_b200 vpb_p1 vpb_p2 vpb_p2l2 vpb_p1t vpic_b200; do
  VPIC_B200_LIB=$PWD/vpic_b200/lib$v.so timeout 600 python bench.py $B > $O/r2r_$v.json 2> $O/r2r_$v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2r_$v.json').read().strip().splitlines()[-1])
r=d['roofline']; print('$v', 'Gpush/s %.2f ms/step %.3f'%(d['value']/1e9, d['ms_per_step']), 'frac', r['frac'], r['ms_per_launch_by_species'], 'sort %.2f'%d['sort_p']['ms_per_call'])
PY
done
