#!/bin/bash
# Round 2, GPU session c: LPI 2-rank drop-in log, lazy-sort tests, regression, bench with lazy sort at several sort intervals
O=gpurun_out; mkdir -p $O
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
mkdir -p /tmp/lpi2 && (cd /tmp/lpi2 && VPIC_B200_DEPOSIT=parity VPIC_B200_COHERENCE=strict VPIC_B200_PUT_TIMEOUT_S=60 VPIC_B200_REPORT=1 timeout 600 python $OLDPWD/oracle/mprun.py -n 2 --tag-output -- $OLDPWD/oracle/_ref/deckrun_lpi_small_b200 --tpp 1 > $OLDPWD/$O/r2c_lpi2.log 2>&1; echo "lpi2 rc=$?")
grep -n "rror\|vpic_b200\|Abort\|abort" $O/r2c_lpi2.log | head -20
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/r2c_parity.log 2>&1; echo "parity rc=$?"; tail -n 12 $O/r2c_parity.log
timeout 2400 python -m pytest tests/test_c1_energy_history.py tests/test_c4_full_size_properties.py tests/test_gpu_weibel.py tests/test_multi_rank.py -q -m gpu > $O/r2c_regress.log 2>&1; echo "regress rc=$?"; tail -n 12 $O/r2c_regress.log
B="python bench.py --steps 8 --warmup 3 --no-cpu --no-e2e --no-yee --no-c1 --no-decks"
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('$1', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"; }
VPIC_B200_LAZY_SORT=0 timeout 600 $B 2>$O/r2c_nolazy.err | tail -n 1 | tee $O/r2c_nolazy.json | line nolazy_si8
for si in 8 4,8 3,12 2,8 4,16; do
  timeout 600 $B --sort-interval $si 2>$O/r2c_lazy_si$si.err | tail -n 1 | tee $O/r2c_lazy_si$si.json | line lazy_si$si
done
for si in 8 3,12; do
  VPIC_B200_LIB=$PWD/vpic_b200/libvpb_hr.so timeout 600 $B --sort-interval $si 2>$O/r2c_hr_si$si.err | tail -n 1 | tee $O/r2c_hr_si$si.json | line hr_lazy_si$si
done
timeout 600 $B --steps 12 --warmup 3 --sort-interval 3,12 --breakdown 12 2>$O/r2c_breakdown.err | tail -n 1 > $O/r2c_breakdown.json; python -c "
import json; d=json.load(open('$O/r2c_breakdown.json')); print(d['breakdown_ms'])"
ls -la $O | grep r2c
