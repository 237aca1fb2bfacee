#!/bin/bash
# A/B a variant library: tools/ab.sh path/to/lib.so [more libs]  -> kernel time (plain run) and DRAM bytes (ncu) of k_advance_p
for lib in "$@"; do
  export VPIC_B200_LIB=$PWD/$lib
  VPIC_B200_LIB=$PWD/$lib python bench.py --cells 128x128x128 --steps 16 --warmup 4 --no-cpu --no-e2e --no-yee 2>&1 | tail -n 1 |
    LIBN=$lib python -c "import json,sys,os; d=json.loads(sys.stdin.read()); print(os.environ['LIBN'], 'Gpush/s', round(d['value']/1e9,2), 'ms/step', round(d['ms_per_step'],3), 'advance_p ms', round(d['roofline']['ms_per_launch'],4))"
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_advance_p --launch-skip 8 -c 4 --csv python bench.py --cells 128x128x128 --steps 4 --warmup 4 --no-cpu --no-e2e --no-yee 2>/dev/null | grep '^"' | python -c "
import csv,sys
for r in list(csv.reader(sys.stdin))[1:]: print('   ', r[0], r[-3], r[-1], r[-2])"
done
