#!/bin/bash
# branch-free segment pass: parity + timing
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "lazy or sort" > $O/r2ag_parity.log 2>&1; echo "parity rc=$?"; tail -n 3 $O/r2ag_parity.log | cut -c1-300
timeout 900 python -m pytest tests/test_multi_rank.py -x -q -m gpu -k "resident" > $O/r2ag_mr.log 2>&1; echo "mr rc=$?"; tail -n 3 $O/r2ag_mr.log | cut -c1-300
timeout 600 python -m pytest tests/test_c4_full_size_properties.py -x -q -m gpu > $O/r2ag_c4.log 2>&1; echo "c4 rc=$?"; tail -n 2 $O/r2ag_c4.log | cut -c1-300
B="--steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
timeout 600 python bench.py $B > $O/r2ag_bench.json 2> $O/r2ag_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2ag_bench.json').read().strip().splitlines()[-1])
r=d['roofline']; print('Gpush/s %.2f ms/step %.3f'%(d['value']/1e9, d['ms_per_step']), 'frac', r['frac'], r['ms_per_launch_by_species'], 'sort', d['sort_p']['ms_per_call'])
PY
C128="python bench.py --cells 128x128x128 --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_seg' -c 20 --csv --log-file $O/r2ag_seg_128.csv $C128 > $O/r2ag_ncu.log 2>&1; echo "ncu rc=$?"; grep "k_seg" $O/r2ag_seg_128.csv | awk -F'","' '{print $5, $NF}' | cut -c1-90
