O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/c8_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/c8_pytest.log
tools/gpu_call.sh c8 vpic_b200/libvpic_b200.so
python bench.py --cells 128x128x128 --steps 16 --warmup 4 --no-cpu --no-e2e --no-yee --breakdown 16 --sort-interval 4,16 | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('si=4,16', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'])); print(d['breakdown_ms'])"
