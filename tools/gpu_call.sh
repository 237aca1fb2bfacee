#!/bin/bash
# One GPU session of the round: sanity (pytest -m gpu), A/B of k_advance_p variants, ncu capture.
# usage: tools/gpu_call.sh TAG [lib ...]   (outputs under gpurun_out/)
TAG=$1; shift
O=gpurun_out; mkdir -p $O
B="python bench.py --cells 128x128x128 --steps 16 --warmup 4 --no-cpu --no-e2e --no-yee"
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('$1', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"; }
for lib in "$@"; do
  name=$(basename $lib .so)
  VPIC_B200_LIB=$PWD/$lib $B 2>$O/${TAG}_$name.err | tail -n 1 | tee $O/${TAG}_$name.json | line $name
done
