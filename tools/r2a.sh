#!/bin/bash
# Round 2, GPU session a: no-alloc test alone, full GPU suite, full-size baseline + L2-hint A/B, full-size ncu counters of k_advance_p
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > $O/r2a_gpu.txt
free -g | head -2 >> $O/r2a_gpu.txt; nproc >> $O/r2a_gpu.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/r2a_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
tail -n 3 $O/r2a_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu > $O/r2a_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 5 $O/r2a_pytest.log
B="python bench.py --steps 8 --warmup 3 --no-cpu --no-e2e --no-yee"
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['ms_per_launch_by_species']
print('$1', 'Gpush/s %.2f ms/step %.3f advance_p %.4f ion %.4f electron %.4f frac %.3f sort %.3f' % (d['value']/1e9, d['ms_per_step'], r['ms_per_launch'], s['ion'], s['electron'], r['frac'], d['sort_p']['ms_per_call']))"; }
for lib in vpic_b200/libvpic_b200.so vpic_b200/libvpb_hs.so vpic_b200/libvpb_hr.so vpic_b200/libvpb_hsr.so; do
  name=$(basename $lib .so)
  VPIC_B200_LIB=$PWD/$lib timeout 600 $B 2>$O/r2a_$name.err | tail -n 1 | tee $O/r2a_$name.json | line $name
done
# full-size counters: launches 8..11 = steps 4, 5 of the sort interval that starts with the warm-up's first step
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_op_red.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,l1tex__data_pipe_lsu_wavefronts.sum,smsp__inst_executed.sum,sm__cycles_active.avg,l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum,lts__t_sectors_srcunit_tex_op_red_lookup_miss.sum
for lib in vpic_b200/libvpic_b200.so vpic_b200/libvpb_hsr.so; do
  name=$(basename $lib .so)
  VPIC_B200_LIB=$PWD/$lib timeout 1200 ncu --metrics $M --clock-control none -k regex:k_advance_p --launch-skip 8 -c 4 --csv --log-file $O/r2a_fullsize_$name.csv python bench.py --steps 4 --warmup 3 --no-cpu --no-e2e --no-yee > $O/r2a_ncu_$name.log 2>&1; echo "ncu $name rc=$?"
done
ls -la $O | grep r2a
