O=gpurun_out; mkdir -p $O
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']; n=d['config']['particles_per_gpu']/2
print('$1', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f | ps/push ion %.2f electron %.2f sort %.2f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call'], s['ion']*1e9/n, s['electron']*1e9/n, d['sort_p']['ms_per_call']*1e9/n))"; }
B="python bench.py --steps 16 --warmup 3 --no-cpu --no-e2e --no-yee"
$B --cells 256x128x128 2>/dev/null | tail -n 1 | line 256x128x128
$B --cells 256x256x256 2>/dev/null | tail -n 1 | line 256x256x256
$B --cells 256x256x256 --ppc 64 2>/dev/null | tail -n 1 | line 256x256x256_ppc64
$B --cells 512x256x256 --ppc 16 2>/dev/null | tail -n 1 | line 512x256x256_ppc16
$B --cells 1024x512x8 2>/dev/null | tail -n 1 | line 1024x512x8
