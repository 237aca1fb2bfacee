#!/bin/bash
# Round 2, GPU session d: LPI N-rank drop-ins, lazy + segment sort tests, bench at several sort intervals, prefetch variant
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/r2d_parity.log 2>&1; echo "parity rc=$?"; tail -n 12 $O/r2d_parity.log
timeout 2400 python -m pytest tests/test_gpu_dropin.py -q -m gpu -k "n_ranks or pcomm or restore" > $O/r2d_dropin.log 2>&1; echo "dropin rc=$?"; tail -n 8 $O/r2d_dropin.log
timeout 2400 python -m pytest tests/test_c1_energy_history.py tests/test_c4_full_size_properties.py tests/test_gpu_weibel.py -q -m gpu > $O/r2d_regress.log 2>&1; echo "regress rc=$?"; tail -n 6 $O/r2d_regress.log
B="python bench.py --steps 12 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('$1', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"; }
VPIC_B200_SEGMENT_SORT=0 timeout 600 $B --sort-interval 3,12 2>$O/r2d_noseg.err | tail -n 1 | tee $O/r2d_noseg.json | line noseg_si3,12
for si in 8 3,12 2,12 2,8 1,8; do
  timeout 600 $B --sort-interval $si 2>$O/r2d_seg_si$si.err | tail -n 1 | tee $O/r2d_seg_si$si.json | line seg_si$si
done
for lib in hr hrp; do for si in 3,12 2,12; do
  VPIC_B200_LIB=$PWD/vpic_b200/libvpb_$lib.so timeout 600 $B --sort-interval $si 2>$O/r2d_${lib}_si$si.err | tail -n 1 | tee $O/r2d_${lib}_si$si.json | line ${lib}_si$si
done; done
timeout 600 $B --sort-interval 2,12 --breakdown 12 2>$O/r2d_breakdown.err | tail -n 1 > $O/r2d_breakdown.json; python -c "
import json; d=json.load(open('$O/r2d_breakdown.json')); print(d['breakdown_ms'])"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_seg|k_scan|k_sort|k_extract' -c 60 --csv --log-file $O/r2d_sort_launches.csv python bench.py --cells 128x128x128 --steps 6 --warmup 3 --sort-interval 2,4 --no-cpu --no-e2e --no-yee --no-c1 --no-decks > $O/r2d_ncu_sort.log 2>&1; echo "ncu sort rc=$?"
timeout 300 tools/pciebench > $O/r2d_pcie.txt 2>&1; cat $O/r2d_pcie.txt
ls -la $O | grep r2d
