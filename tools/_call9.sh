O=gpurun_out; mkdir -p $O
( time python bench.py > $O/c9_full.json 2> $O/c9_full.err ) 2>&1 | grep real; tail -n 1 $O/c9_full.json | cut -c1-1500
for si in 6,24 5,20; do
python bench.py --steps 24 --warmup 3 --no-cpu --no-e2e --no-yee --sort-interval $si 2>/dev/null | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('full si=$si', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"
done
python tools/e2e_breakdown.py 2>&1 | tail -n 18
