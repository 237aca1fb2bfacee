#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sort or lazy or parity_mode" > $O/r2n_parity.log 2>&1; echo "parity rc=$?"; tail -n 4 $O/r2n_parity.log
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('$1', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"; }
B="python bench.py --steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for si in 3,12 2,12 2,8; do
  timeout 600 $B --sort-interval $si 2>>$O/r2n.err | tail -n 1 | line runs_full_si$si | tee -a $O/r2n_ab.txt
done
VPIC_B200_LAZY_SORT=0 timeout 600 $B --sort-interval 8 2>>$O/r2n.err | tail -n 1 | line runs_full_nolazy_si8 | tee -a $O/r2n_ab.txt
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_sort|k_hist|k_scan' --launch-skip 10 -c 12 --csv --log-file $O/r2n_sort_launches.csv python bench.py --cells 128x128x128 --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks > $O/r2n_ncu.log 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(l for l in open("gpurun_out/r2n_sort_launches.csv") if l.startswith('"')))
ix={h:i for i,h in enumerate(rows[0])}
for r in rows[1:]: print(r[ix["Kernel Name"]].split("(")[0][:30], r[ix["Metric Value"]], r[ix["Metric Unit"]])
PY
