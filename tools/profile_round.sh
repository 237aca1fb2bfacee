#!/bin/bash
# Round profile set (B200_PROFILING.md recipe): the plain run first, then the
# launch list and one --set full capture of the hot kernels of the same command.
# usage: tools/profile_round.sh TAG      (outputs under gpurun_out/)
TAG=${1:-r1}
CMD="python bench.py --cells 128x128x128 --steps 8 --warmup 4 --no-cpu --no-e2e"
O=gpurun_out
$CMD > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err || { echo "plain run failed"; tail -n 5 $O/${TAG}_bench.err; exit 1; }
tail -n 1 $O/${TAG}_bench.json | cut -c1-400
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${TAG}_launches.csv $CMD --no-yee > $O/${TAG}_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_advance_p|k_sort|k_load_interp|k_unload|k_scan' --launch-skip 60 -c 24 -o $O/${TAG}_particles -f $CMD --no-yee > $O/${TAG}_ncu_p.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_advance_b|k_advance_e' --launch-skip 40 -c 6 -o $O/${TAG}_yee -f python bench.py --cells 32x32x32 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/${TAG}_ncu_y.log 2>&1
ls -la $O | grep $TAG
