O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/c2_pytest_main.log 2>&1; echo "pytest main rc=$?"; tail -n 2 $O/c2_pytest_main.log
VPIC_B200_LIB=$PWD/vpic_b200/libvpb_sq.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/c2_pytest_sq.log 2>&1; echo "pytest sq rc=$?"; tail -n 2 $O/c2_pytest_sq.log
tools/gpu_call.sh c2 vpic_b200/libvpic_b200.so vpic_b200/libvpb_sq.so
B="python bench.py --cells 128x128x128 --steps 32 --warmup 4 --no-cpu --no-e2e --no-yee"
for si in 4,16 4,32 6,24 8,32 5,20; do
  for lib in vpic_b200/libvpic_b200.so vpic_b200/libvpb_sq.so; do
  VPIC_B200_LIB=$PWD/$lib $B --sort-interval $si 2>/dev/null | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('$lib si=$si', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"
  done
done
