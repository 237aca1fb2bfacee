O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/c17_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 $O/c17_pytest.log
python bench.py --cells 128x128x128 --steps 16 --warmup 4 --no-cpu --no-e2e 2>/dev/null | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('128^3 Gpush/s %.2f ms/step %.3f advance_p %.4f launches/step %.1f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], d['gpu_launches']/d['steps']))
print({k:(round(v['ms'],4),round(v['frac_of_hbm_peak'],3)) for k,v in d['yee'].items() if isinstance(v,dict)})"
