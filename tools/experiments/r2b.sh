#!/bin/bash
# Round 2, GPU session b: new multi-rank / drop-in tests on one GPU, advance_p variant A/B at full size, 128^3 source-level capture
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/r2b_parity.log 2>&1; echo "parity rc=$?"; tail -n 3 $O/r2b_parity.log
timeout 2400 python -m pytest tests/test_multi_rank.py -q -m gpu > $O/r2b_multirank.log 2>&1; echo "multirank rc=$?"; tail -n 15 $O/r2b_multirank.log
timeout 3000 python -m pytest tests/test_gpu_dropin.py -q -m gpu > $O/r2b_dropin.log 2>&1; echo "dropin rc=$?"; tail -n 25 $O/r2b_dropin.log
B="python bench.py --steps 8 --warmup 3 --no-cpu --no-e2e --no-yee"
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('$1', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"; }
for lib in vpic_b200/libvpic_b200.so vpic_b200/libvpb_hr.so vpic_b200/libvpb_c.so vpic_b200/libvpb_hrc.so vpic_b200/libvpb_hsrc.so; do
  name=$(basename $lib .so)
  VPIC_B200_LIB=$PWD/$lib timeout 600 $B 2>$O/r2b_$name.err | tail -n 1 | tee $O/r2b_$name.json | line $name
done
for si in 4,16 6,24 4,8; do
  VPIC_B200_LIB=$PWD/vpic_b200/libvpb_hrc.so timeout 600 $B --sort-interval $si 2>$O/r2b_hrc_si$si.err | tail -n 1 | tee $O/r2b_hrc_si$si.json | line hrc_si$si
done
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_op_red.sum,l1tex__data_pipe_lsu_wavefronts.sum,smsp__inst_executed.sum,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum,lts__t_sectors_srcunit_tex_op_red_lookup_miss.sum
VPIC_B200_LIB=$PWD/vpic_b200/libvpb_hrc.so timeout 1200 ncu --metrics $M --clock-control none -k regex:k_advance_p --launch-skip 8 -c 2 --csv --log-file $O/r2b_fullsize_hrc.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-yee > $O/r2b_ncu_hrc.log 2>&1; echo "ncu hrc rc=$?"
VPIC_B200_LIB=$PWD/vpic_b200/libvpb_hrc.so timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_advance_p --launch-skip 8 -c 2 -o $O/r2b_ap128 -f python bench.py --cells 128x128x128 --steps 2 --warmup 3 --no-cpu --no-e2e --no-yee > $O/r2b_ncu_ap128.log 2>&1; echo "ncu 128 rc=$?"
ls -la $O | grep r2b
