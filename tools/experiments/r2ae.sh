#!/bin/bash
O=$PWD/gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q > $O/r2ae_parity.log 2>&1; echo "parity rc=$?"; tail -n 2 $O/r2ae_parity.log | cut -c1-300
timeout 2400 python -m pytest tests/test_gpu_dropin.py -x -q > $O/r2ae_dropin.log 2>&1; echo "dropin rc=$?"; tail -n 4 $O/r2ae_dropin.log | cut -c1-300
