#!/bin/bash
# segment sort: parity + A/B at full size + kernel times at 128^3
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > $O/r2t_parity.log 2>&1; echo "parity rc=$?"; tail -n 15 $O/r2t_parity.log | cut -c1-300
timeout 1500 python -m pytest tests/test_multi_rank.py -x -q -m gpu -k "resident" > $O/r2t_mr.log 2>&1; echo "mr rc=$?"; tail -n 12 $O/r2t_mr.log | cut -c1-400
timeout 900 python -m pytest tests/test_c4_full_size_properties.py -x -q -m gpu > $O/r2t_c4.log 2>&1; echo "c4 rc=$?"; tail -n 8 $O/r2t_c4.log | cut -c1-300
B="--steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for v in 1 0 1 0; do
  VPIC_B200_SEG_SORT=$v timeout 600 python bench.py $B > $O/r2t_seg$v.json 2> $O/r2t_seg$v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2t_seg$v.json').read().strip().splitlines()[-1])
r=d['roofline']; print('seg=$v', 'Gpush/s %.2f ms/step %.3f'%(d['value']/1e9, d['ms_per_step']), 'frac', r['frac'], r['ms_per_launch_by_species'], 'sort %.2f'%d['sort_p']['ms_per_call'])
PY
done
C128="python bench.py --cells 128x128x128 --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_sort|k_hist|k_scan|k_seg' -c 80 --csv --log-file $O/r2t_sort_launches_128.csv $C128 > $O/r2t_ncu.log 2>&1; echo "ncu rc=$?"
