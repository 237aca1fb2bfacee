#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "carried_out or tracked or langevin" > $O/r2j_parity.log 2>&1; echo "parity rc=$?"; tail -n 15 $O/r2j_parity.log
timeout 2400 python -m pytest tests/test_gpu_weibel.py tests/test_gpu_dropin.py tests/test_zz_no_alloc.py -x -q -m gpu > $O/r2j_rest.log 2>&1; echo "rest rc=$?"; tail -n 15 $O/r2j_rest.log
