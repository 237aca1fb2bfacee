#!/bin/bash
# Round 2, GPU session h: the whole GPU suite as the driver runs it, then smoke()
O=gpurun_out; mkdir -p $O
( time timeout 3000 python -m pytest tests -x -q -m gpu ) > $O/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 12 $O/r2h_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2h_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/r2h_smoke.log
