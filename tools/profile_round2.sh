#!/bin/bash
# Round-2 profile set (B200_PROFILING.md recipe): plain default bench first, then the launch list of the same command,
# the DRAM counters of k_advance_p at the benchmarked size, and --set full captures at 128^3.
# usage: tools/profile_round2.sh TAG      (outputs under gpurun_out/)
TAG=${1:-r2k}
O=gpurun_out; mkdir -p $O
timeout 1500 python bench.py > $O/${TAG}_bench_1gpu.json 2> $O/${TAG}_bench_1gpu.err || { echo "plain run failed"; tail -n 5 $O/${TAG}_bench_1gpu.err; exit 1; }
python - <<PY
import json
d=json.loads(open('$O/${TAG}_bench_1gpu.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches')}); r=d['roofline']; print('roofline', r['frac'], r['ms_per_launch_by_species'], r['ms_by_steps_since_sort'], 'share', r['share_of_step']); print('sort', d['sort_p'])
print('e2e', {k: d['e2e'][k] for k in d['e2e'] if k not in ('workload','energies_after')})
for k in ('harris_c3','lpi_c5'):
    x=d['deck_e2e'][k]; print(k, {a: x[a] for a in x if a.startswith('speedup')}, 'gpu', x['gpu'].get('steady_ms_per_step'), 'cpu', x['cpu'].get('steady_ms_per_step'))
print('c1', d['c1']['pushes_per_s'], d['c1']['ms_per_step_device'], d['c1']['gpu_launches_per_step']); print('cpu', d['cpu_baseline']['value']); print('yee', d['yee']['field_step']); print('clocks', d['clocks'])
PY
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/${TAG}_launches.csv $CMD > $O/${TAG}_ncu_list.log 2>&1; echo "launch list rc=$?"
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,l1tex__data_pipe_lsu_wavefronts.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
ncu --metrics $M --clock-control none -k regex:k_advance_p --launch-skip 6 -c 6 --csv --log-file $O/${TAG}_fullsize_advance_p.csv $CMD > $O/${TAG}_ncu_fs.log 2>&1; echo "full-size counters rc=$?"
C128="python bench.py --cells 128x128x128 --steps 3 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
ncu --set full --clock-control none --import-source on -k regex:'k_advance_p' --launch-skip 6 -c 4 -o $O/${TAG}_ap128 -f $C128 > $O/${TAG}_ncu_ap128.log 2>&1; echo "ncu advance_p 128 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_sort|k_hist|k_scan|k_seg' --launch-skip 12 -c 12 -o $O/${TAG}_sort128 -f $C128 > $O/${TAG}_ncu_sort128.log 2>&1; echo "ncu sort 128 rc=$?"
ls -la $O | grep ${TAG}
