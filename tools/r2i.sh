#!/bin/bash
O=gpurun_out; mkdir -p $O
( time timeout 3000 python -m pytest tests -x -q -m gpu ) > $O/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 12 $O/r2i_pytest.log
