#!/bin/bash
# what the driver runs at round end, on one GPU: the GPU test suite, smoke(), the default bench line
O=gpurun_out; mkdir -p $O
timeout 3000 python -m pytest tests -x -q -m gpu > $O/final_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -n 4 $O/final_gpu_tests.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 $O/final_smoke.log | cut -c1-200
timeout 1500 python bench.py > $O/final_bench_1gpu.json 2> $O/final_bench_1gpu.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/final_bench_1gpu.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches')}); r=d['roofline']; print('roofline', r['frac'], r['traffic'], r['ms_per_launch_by_species']); print('sort', d['sort_p']['ms_per_call'], d['sort_p']['effective'])
print('e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'strict', d['e2e']['strict']['value'])
for k in ('harris_c3','lpi_c5'):
    x=d['deck_e2e'][k]; print(k, {a: x[a] for a in x if a.startswith('speedup')}, 'walls', x['gpu']['wall_s'], x['cpu']['wall_s'], x.get('gpu_unmodified_deck',{}).get('wall_s'))
print('c1', d['c1']['pushes_per_s'], 'cpu', d['cpu_baseline']['value'], 'yee', d['yee']['field_step']['cell_updates_per_s'], 'clocks', d['clocks'])
PY
