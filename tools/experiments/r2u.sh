#!/bin/bash
# segment sort, second version: parity + A/B at full size + kernel times at 128^3 and at full size
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "lazy or sort" > $O/r2u_parity.log 2>&1; echo "parity rc=$?"; tail -n 5 $O/r2u_parity.log | cut -c1-300
timeout 1500 python -m pytest tests/test_multi_rank.py -x -q -m gpu -k "resident" > $O/r2u_mr.log 2>&1; echo "mr rc=$?"; tail -n 5 $O/r2u_mr.log | cut -c1-400
B="--steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for v in 1 0 1 0; do
  VPIC_B200_SEG_SORT=$v timeout 600 python bench.py $B > $O/r2u_seg$v.json 2> $O/r2u_seg$v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2u_seg$v.json').read().strip().splitlines()[-1])
r=d['roofline']; print('seg=$v', 'Gpush/s %.2f ms/step %.3f'%(d['value']/1e9, d['ms_per_step']), 'frac', r['frac'], r['ms_per_launch_by_species'], 'sort %.2f'%d['sort_p']['ms_per_call'])
PY
done
C128="python bench.py --cells 128x128x128 --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_sort|k_hist|k_scan|k_seg' -c 80 --csv --log-file $O/r2u_sort_launches_128.csv $C128 > $O/r2u_ncu.log 2>&1; echo "ncu rc=$?"
CF="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_sort|k_hist|k_scan|k_seg' -c 80 --csv --log-file $O/r2u_sort_launches_full.csv $CF > $O/r2u_ncu_full.log 2>&1; echo "ncu full rc=$?"
