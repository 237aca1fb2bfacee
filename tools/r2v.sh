#!/bin/bash
# segment sort with aggregated stray atomics: parity, cadence A/B, kernel times at full size
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "lazy or sort" > $O/r2v_parity.log 2>&1; echo "parity rc=$?"; tail -n 5 $O/r2v_parity.log | cut -c1-300
timeout 1500 python -m pytest tests/test_multi_rank.py -x -q -m gpu -k "resident" > $O/r2v_mr.log 2>&1; echo "mr rc=$?"; tail -n 5 $O/r2v_mr.log | cut -c1-400
B="--steps 24 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for si in 3,12 3,8 3,6 2,8 3,12; do
  timeout 600 python bench.py $B --sort-interval $si > $O/r2v_si$si.json 2> $O/r2v_si$si.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2v_si$si.json').read().strip().splitlines()[-1])
r=d['roofline']; print('si=$si', 'Gpush/s %.2f ms/step %.3f'%(d['value']/1e9, d['ms_per_step']), 'frac', r['frac'], r['ms_per_launch_by_species'], 'sort', d['sort_p']['ms_per_call'], d['sort_p'].get('effective'))
PY
done
CF="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_seg' -c 20 --csv --log-file $O/r2v_sort_launches_full.csv $CF > $O/r2v_ncu_full.log 2>&1; echo "ncu full rc=$?"
