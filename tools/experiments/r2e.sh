#!/bin/bash
# Round 2, GPU session e: TRACKED + restore tests, advance_p variant A/B at 128^3 (interleaved, twice), full default bench
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tracked or lazy" > $O/r2e_parity.log 2>&1; echo "parity rc=$?"; tail -n 12 $O/r2e_parity.log
timeout 1500 python -m pytest tests/test_gpu_dropin.py -q -m gpu -k "restore" > $O/r2e_dropin.log 2>&1; echo "dropin rc=$?"; tail -n 5 $O/r2e_dropin.log
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('$1', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"; }
B="python bench.py --cells 128x128x128 --steps 24 --warmup 4 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for rep in 1 2; do for lib in vpic_b200 vpb_hr vpb_hrp vpb_hrr vpb_r; do for si in 8 3,12; do
  VPIC_B200_LIB=$PWD/vpic_b200/lib$lib.so timeout 300 $B --sort-interval $si 2>>$O/r2e_ab.err | tail -n 1 | line ${lib}_si${si}_rep$rep | tee -a $O/r2e_ab.txt
done; done; done
B="python bench.py --steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
for lib in vpb_hr vpb_hrr; do
  VPIC_B200_LIB=$PWD/vpic_b200/lib$lib.so timeout 600 $B --sort-interval 3,12 2>>$O/r2e_ab.err | tail -n 1 | line full_${lib}_si3,12 | tee -a $O/r2e_ab.txt
done
timeout 1500 python bench.py > $O/r2e_bench_default.json 2> $O/r2e_bench_default.err; echo "default bench rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2e_bench_default.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches')}); print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch_by_species'])
print('e2e', {k: d['e2e'][k] for k in d['e2e'] if k not in ('workload','energies_after')})
print('c1', d.get('c1')); print('deck_e2e', json.dumps(d.get('deck_e2e'))[:1500]); print('cpu', d.get('cpu_baseline')); print('yee', d.get('yee',{}).get('field_step'))
PY
ls -la $O | grep r2e
