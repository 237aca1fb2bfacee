#!/bin/bash
# Round 2, GPU session f: TRACKED test with output, deck stage tables, full default bench
O=gpurun_out; mkdir -p $O
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tracked" -s -p no:faulthandler > $O/r2f_tracked.log 2>&1; echo "tracked rc=$?"; tail -n 15 $O/r2f_tracked.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tracked" > $O/r2f_tracked2.log 2>&1; echo "tracked (default plugins) rc=$?"; tail -n 5 $O/r2f_tracked2.log
for deck in harris_c3 lpi_c5; do
  mkdir -p /tmp/$deck && (cd /tmp/$deck && VPIC_B200_COHERENCE=resident VPIC_B200_DEPOSIT=fast VPIC_B200_REPORT=1 timeout 600 $OLDPWD/oracle/_ref/deckrun_${deck}_res_b200 --tpp 1 > $OLDPWD/$O/r2f_$deck.log 2>&1; echo "$deck rc=$?")
  grep -n "|" $O/r2f_$deck.log | tail -45 | cut -c1-150
  grep "vpic_b200\]" $O/r2f_$deck.log | tail -2
done
timeout 1500 python bench.py > $O/r2f_bench_default.json 2> $O/r2f_bench_default.err; echo "default bench rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2f_bench_default.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches')}); print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch_by_species'], 'sort', d['sort_p'])
print('e2e', {k: d['e2e'][k] for k in d['e2e'] if k not in ('workload','energies_after')})
for k in ('harris_c3','lpi_c5'):
    x=d['deck_e2e'][k]; print(k, {a: x[a] for a in x if a.startswith('speedup')}, 'gpu', x['gpu'].get('steady_ms_per_step'), x['gpu'].get('stages_s'), 'cpu', x['cpu'].get('steady_ms_per_step'))
PY
ls -la $O | grep r2f
