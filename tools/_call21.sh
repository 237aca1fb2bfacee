O=gpurun_out; mkdir -p $O
( time python bench.py > $O/c21_full.json 2> $O/c21_full.err ) 2>&1 | grep real; tail -n 1 $O/c21_full.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.2f G ms/step %.3f frac %.4f adv %.3f' % (d['value']/1e9, d['ms_per_step'], r['frac'], r['ms_per_launch']), r['ms_per_launch_by_species'])
print('e2e', {k:v for k,v in d['e2e'].items() if k not in ('workload','energies_after')})
print('cpu', {k:v for k,v in d['cpu_baseline'].items()})
print('yee', {k:(round(v['ms'],4),round(v['frac_of_hbm_peak'],3)) for k,v in d['yee'].items() if isinstance(v,dict)}, d['clocks'], d['gpu_launches'])"
( time python bench.py --impl reference --steps 10 --warmup 3 > $O/c21_ref.json 2> $O/c21_ref.err ) 2>&1 | grep real; tail -n 1 $O/c21_ref.json | cut -c1-900
